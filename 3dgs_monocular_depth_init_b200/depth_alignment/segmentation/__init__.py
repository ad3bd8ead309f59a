"""Depth-map segmenters for region-wise alignment (reference: depth_alignment/segmentation/).

The SAM and SLIC segmenters and the RAG region merging of the reference need segment_anything /
scikit-image and are outside the B200 path.  What the pipeline needs from them
is an integer region map, so a segmenter here is any callable with the reference's signature

    fn(predicted_depth: PredictedDepth, checkpoint_dir: Path, config: DepthSegmentationConfig)
        -> int tensor [H, W] of region ids (already merged; ids 0 .. R-1)

registered under the name the config selects (`mdi.alignment.segmenter = "slic" | "sam"`)."""
from typing import Callable, Dict

_SEGMENTERS: Dict[str, Callable] = {}


def register_segmenter(name: str, fn) -> None:
    """Plug a segmentation function in for `config.mdi.alignment.segmenter == name` (None removes it)."""
    if fn is None:
        _SEGMENTERS.pop(name, None)
    else:
        _SEGMENTERS[name] = fn


def get_segmenter(name: str) -> Callable:
    fn = _SEGMENTERS.get(getattr(name, "value", name))
    if fn is None:
        raise NotImplementedError(
            f"Depth segmentation '{getattr(name, 'value', name)}' is out of scope of the B200 path (it needs "
            "segment_anything / scikit-image): register a function returning the integer region map with "
            "depth_alignment.segmentation.register_segmenter()")
    return fn
