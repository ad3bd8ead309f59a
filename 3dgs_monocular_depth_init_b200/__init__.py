"""B200-native monocular-depth point-cloud initialisation (drop-in for the hot path of
deivse/3dgs_monocular_depth_init).  Import with importlib (the name starts with a digit):

    mdi = importlib.import_module("3dgs_monocular_depth_init_b200")
"""
from . import _lib
from ._lib import MdiError
from .build import build

__all__ = ["build", "MdiError", "_lib"]
