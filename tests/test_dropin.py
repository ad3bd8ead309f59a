"""The drop-in boundary on the CPU (no kernel launches): the `gs_init_compare` alias package, the
overlay installer, the predicted-depth cache reader, PLY export, and config-field handling.
Reference: runner.py:42,77-78 (import + call site), monocular_depth_init.py:32-87,198-213."""
import subprocess
import sys
import types
from pathlib import Path

import numpy as np
import pytest
import torch

from helpers import ROOT, mod


def _run_py(code: str):
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=str(ROOT))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r.stdout


def test_alias_package_resolves_reference_import_paths():
    """Own process: the CPU suite also imports the real reference as `gs_init_compare` (ref_import)."""
    out = _run_py(
        "import importlib, sys\n"
        "sys.path.insert(0, '.')\n"
        "from gs_init_compare.monocular_depth_init import pts_and_rgb_from_monocular_depth as f\n"   # runner.py:42
        "from gs_init_compare.config import Config\n"
        "from gs_init_compare.depth_prediction.points_from_depth import get_pts_from_depth\n"
        "from gs_init_compare.depth_alignment.exceptions import LowDepthAlignmentConfidenceError as E\n"
        "from gs_init_compare.depth_prediction.predictors.depth_predictor_interface import PredictedDepth as PD\n"
        "from gs_init_compare.utils.runner_utils import knn\n"
        "from gs_init_compare.utils.point_cloud_export import export_point_cloud_to_ply\n"
        "impl = importlib.import_module('3dgs_monocular_depth_init_b200.monocular_depth_init')\n"
        "assert f is impl.pts_and_rgb_from_monocular_depth\n"
        "exc = importlib.import_module('3dgs_monocular_depth_init_b200.depth_alignment.exceptions')\n"
        "assert E is exc.LowDepthAlignmentConfidenceError\n"
        "import inspect; assert list(inspect.signature(f).parameters)[:3] == ['config', 'parser', 'device']\n"
        "c = Config(); assert c.mdi.alignment.aligner.get_implementation().__name__ == 'DepthAlignmentRansac'\n"
        "try:\n    import gs_init_compare.runner\n    raise SystemExit('runner must not resolve')\n"
        "except ModuleNotFoundError: pass\n"
        "print('alias ok')\n")
    assert "alias ok" in out


def test_overlay_install_rebinds_an_installed_reference():
    """dropin.install() on a stand-in package tree shaped like the reference."""
    pkg = "fake_ref_pkg"
    tree = ["", "monocular_depth_init", "depth_prediction", "depth_prediction.points_from_depth", "runner",
            "depth_alignment", "depth_alignment.pipeline", "utils", "utils.runner_utils"]
    try:
        for rel in tree:
            name = pkg + ("." + rel if rel else "")
            m = types.ModuleType(name)
            m.__path__ = []
            sys.modules[name] = m
        sentinel = lambda *a, **k: "reference"
        sys.modules[pkg + ".monocular_depth_init"].pts_and_rgb_from_monocular_depth = sentinel
        sys.modules[pkg + ".runner"].pts_and_rgb_from_monocular_depth = sentinel   # `from ... import` copy
        sys.modules[pkg + ".utils.runner_utils"].knn = sentinel
        done = mod("dropin").install(pkg)
        ours = mod("monocular_depth_init").pts_and_rgb_from_monocular_depth
        assert sys.modules[pkg + ".monocular_depth_init"].pts_and_rgb_from_monocular_depth is ours
        assert sys.modules[pkg + ".runner"].pts_and_rgb_from_monocular_depth is ours
        assert sys.modules[pkg + ".utils.runner_utils"].knn is mod("knn").knn
        assert f"{pkg}.runner.pts_and_rgb_from_monocular_depth" in done
    finally:
        for k in [k for k in sys.modules if k == pkg or k.startswith(pkg + ".")]:
            del sys.modules[k]


def test_model_name_and_cache_path_follow_the_reference():
    cfgm, dc, pc = mod("config"), mod("depth_prediction.depth_cache"), mod("depth_prediction.configs")
    c = cfgm.Config()
    assert dc.model_name(c) == "Metric3d_vitl"                                   # metric3d.py:26
    c.mdi.metric3d.backbone = pc.Metric3dBackbone.vitg
    assert dc.model_name(c) == "Metric3d_vitg"
    assert str(dc.cache_path(c, "garden", "DSC07956")) == "__mono_depth_cache__/Metric3d_vitg/garden/DSC07956.pth"
    c.mdi.predictor = "moge"; assert dc.model_name(c) == "MoGe_vitl"            # moge.py:15
    c.mdi.predictor = "unidepth"; assert dc.model_name(c) == "UniDepth_vitl"    # unidepth.py:38
    c.mdi.predictor = "depth_pro"; assert dc.model_name(c) == "AppleDepthPro"   # apple_depth_pro.py:54
    c.mdi.predictor = "depth_anything_v2"
    assert dc.model_name(c) == "DepthAnythingV2_vitl_metric_indoor"             # depth_anything_v2.py:64
    c.mdi.depthanything.metric = False
    assert dc.model_name(c) == "DepthAnythingV2_vitl_relative"                  # :80
    c.mdi.predictor = None
    with pytest.raises(ValueError, match="No depth predictor"):                  # monocular_depth_init.py:33-34
        dc.model_name(c)
    c.mdi.predictor = "midas"
    with pytest.raises(ValueError, match="Unsupported monodepth model"):         # :57
        dc.model_name(c)


def test_cached_depth_source_hit_miss_ignore_and_registered_predictor(tmp_path):
    cfgm, dc = mod("config"), mod("depth_prediction.depth_cache")
    dpi, types_m = mod("depth_prediction.predictors.depth_predictor_interface"), mod("types")
    c = cfgm.Config()
    c.mdi.cache_dir = str(tmp_path / "cache")
    img = types_m.InputImage(data=torch.full((4, 5, 3), 200.0), name="im0", cam2world=torch.eye(4), K=torch.eye(3))
    src = dc.CachedDepthSource(c, "scene", "cpu")
    with pytest.raises(dc.DepthCacheMiss, match="im0.pth"):
        src(img, {}, "cpu")
    p = dc.cache_path(c, "scene", "im0")
    p.parent.mkdir(parents=True)
    want = dpi.PredictedDepth(torch.rand(4, 5), torch.ones(4, 5, dtype=torch.bool))
    torch.save(want, p)
    got = src(img, {}, "cpu")
    assert torch.equal(got.depth, want.depth) and torch.equal(got.mask, want.mask)
    c.mdi.ignore_cache = True                      # the reference then always re-predicts: no predictor here
    with pytest.raises(dc.DepthCacheMiss, match="ignore_cache"):
        src(img, {}, "cpu")

    class Fake(dpi.DepthPredictor):
        calls = 0

        def __init__(self, config, device):
            pass

        @property
        def name(self):
            return "Metric3d_vitl"

        def predict_depth(self, im, intr):
            Fake.calls += 1
            assert float(im.max()) <= 1.0 and abs(intr.fx - 1.0) < 1e-6   # image / 255 (monocular_depth_init.py:128)
            return dpi.PredictedDepth(torch.full((4, 5), 2.0), torch.ones(4, 5, dtype=torch.bool))

    dc.register_predictor("metric3d", Fake)
    try:
        src2 = dc.CachedDepthSource(c, "scene", "cpu")
        assert float(src2(img, {}, "cpu").depth[0, 0]) == 2.0 and Fake.calls == 1
        c.mdi.ignore_cache = False                 # the prediction was saved: now a cache hit
        assert float(src2(img, {}, "cpu").depth[0, 0]) == 2.0 and Fake.calls == 1
    finally:
        dc.register_predictor("metric3d", None)


def test_cache_written_by_the_reference_loads(tmp_path):
    """A file pickled under the reference's module path (what its torch.save produced) loads into
    this package's PredictedDepth with torch.load's default weights_only=True."""
    dpi, dc = mod("depth_prediction.predictors.depth_predictor_interface"), mod("depth_prediction.depth_cache")
    ref_mod = dpi.REFERENCE_QUALNAME.rsplit(".", 1)[0]
    fake = types.ModuleType(ref_mod)

    from dataclasses import dataclass
    from typing import Optional

    @dataclass
    class PredictedDepth:   # the reference's class as pickle sees it: module path + field dict
        depth: torch.Tensor
        mask: torch.Tensor
        depth_confidence: Optional[torch.Tensor] = None
        normal: Optional[torch.Tensor] = None
        normal_confidence: Optional[torch.Tensor] = None

    PredictedDepth.__module__ = ref_mod
    PredictedDepth.__qualname__ = "PredictedDepth"
    fake.PredictedDepth = PredictedDepth
    before = set(sys.modules)
    sys.modules[ref_mod] = fake
    try:
        f = tmp_path / "x.pth"
        torch.save(PredictedDepth(torch.arange(6.0).view(2, 3), torch.ones(2, 3, dtype=torch.bool)), f)
    finally:
        # pickle imports the parent packages of the class's module: leave no `gs_init_compare*` entry
        # behind (this process may import the real reference under that name, oracle/ref_import.py)
        for k in set(sys.modules) - before:
            if k == "gs_init_compare" or k.startswith("gs_init_compare."):
                del sys.modules[k]
        sys.meta_path[:] = [m for m in sys.meta_path if type(m).__name__ != "_AliasFinder"]
    got = dc.load_cached_depth(f, "cpu")
    assert isinstance(got, dpi.PredictedDepth) and got.depth.tolist() == [[0, 1, 2], [3, 4, 5]] and got.normal is None


def test_ply_export_layout_and_round_trip(tmp_path):
    pe = mod("utils.point_cloud_export")
    g = np.random.default_rng(0)
    pts = g.normal(size=(100, 3)).astype(np.float32)
    rgb = g.uniform(-0.1, 1.1, size=(100, 3)).astype(np.float32)
    pe.export_point_cloud_to_ply(pts, rgb, tmp_path, "cloud")
    raw = (tmp_path / "cloud.ply").read_bytes()
    head = raw[: raw.index(b"end_header\n") + 11].decode()
    assert head.startswith("ply\nformat binary_little_endian 1.0\n") and "element vertex 100\n" in head
    assert "property double x" in head and "property uchar blue" in head
    assert len(raw) == len(head) + 100 * 27
    p2, c2 = pe.read_ply(tmp_path / "cloud.ply")
    assert np.array_equal(p2, pts.astype(np.float64))
    assert np.array_equal(c2, np.round(np.clip(rgb.astype(np.float64), 0, 1) * 255).astype(np.uint8))
    pe.export_point_cloud_to_ply(pts, None, tmp_path, "nocol", outlier_std_dev=1.0)   # point_cloud_export.py:27-35
    p3, c3 = pe.read_ply(tmp_path / "nocol.ply")
    keep = np.all(np.abs((pts - pts.mean(0)) / pts.std(0)) < 1.0, axis=1)
    assert c3 is None and np.array_equal(p3, pts[keep].astype(np.float64))


def test_config_surface_matches_reference_field_for_field():
    """Every `mdi.*` dotted path of the reference's Config exists here with the same default."""
    import dataclasses
    import ref_import
    if not ref_import.available():
        pytest.skip("reference tree not present (GPU box)")
    R = ref_import.load()
    ours = mod("config").Config().mdi
    theirs = R.Config().mdi

    def walk(a, b, path):
        for f in dataclasses.fields(b):
            assert hasattr(a, f.name), f"missing {path}.{f.name}"
            x, y = getattr(a, f.name), getattr(b, f.name)
            if dataclasses.is_dataclass(y):
                walk(x, y, f"{path}.{f.name}")
            else:
                assert getattr(x, "value", x) == getattr(y, "value", y), f"{path}.{f.name}: {x!r} != {y!r}"

    walk(ours, theirs, "mdi")
