"""`gs_init_compare` import paths for the B200 hot path.

The reference's callers import the path as `gs_init_compare.<module>` (runner.py:42:
`from gs_init_compare.monocular_depth_init import pts_and_rgb_from_monocular_depth`; nerfbaselines
method.py rewrites that same runner).  With this directory on `sys.path` those imports resolve to
the B200 implementation: `gs_init_compare.X.Y` IS the module `3dgs_monocular_depth_init_b200.X.Y`
(the same module object, so classes, enums and exceptions are identical under both names, and
`torch.load` of a depth cache pickled by the reference finds `PredictedDepth`).

Only the modules of the hot path exist; importing anything else of the
reference (`gs_init_compare.runner`, `.trainer`, `.datasets`, predictors ...) raises
ModuleNotFoundError from here.  To overlay an installed reference instead of replacing the package,
see INTEGRATION.md (`dropin.install()`)."""
import importlib
import importlib.abc
import importlib.util
import sys

_IMPL = "3dgs_monocular_depth_init_b200"


class _AliasLoader(importlib.abc.Loader):
    def __init__(self, target):
        self.target = target

    def create_module(self, spec):
        module = importlib.import_module(self.target)   # the implementation module itself
        self._orig = (module.__spec__, getattr(module, "__loader__", None))
        return module

    def exec_module(self, module):
        # the import machinery re-labelled the module with the alias spec: put its own back, so
        # relative imports and reloads inside the implementation keep working
        module.__spec__, module.__loader__ = self._orig


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith(__name__ + "."):
            return None
        impl = _IMPL + fullname[len(__name__):]
        try:
            found = importlib.util.find_spec(impl)
        except ModuleNotFoundError:
            found = None
        if found is None:
            return None
        return importlib.util.spec_from_loader(fullname, _AliasLoader(impl), is_package=found.submodule_search_locations is not None)


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())
