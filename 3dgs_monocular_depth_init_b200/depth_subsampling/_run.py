from .. import _lib as L
from .._single import single_view_inputs
from ..engine import RunOptions, get_engine


def subsample_mask(depth, mask, **opt_kw):
    """Runs only the subsampler predicate (& mask) of mdi_keep_mask on one view."""
    dev = depth.device
    eng = get_engine(dev)
    vin = single_view_inputs(depth, mask, dev)
    opt = RunOptions(aligner=None, use_patch_mask=False, flags=L.KEEP_PRED_MASK | L.KEEP_SUBSAMPLE,
                     want_xyz=False, want_rgb=False, stop_after="mask", **opt_kw)
    res = eng.run(vin, opt)
    if int(res.status[0].item()) & L.STATUS_EMPTY_STATS:
        raise RuntimeError("quantile() input tensor must be non-empty")  # what torch raises
    return res.keep.view(-1).bool()
