"""TEST INFRASTRUCTURE -- builds the reference's OWN native module, from its sources where they lie
under /root/reference/native_modules/subsampling/src (impl.cpp, pointcloud_subsampling.cpp and their
headers, unmodified, never copied), into oracle/_ref/_pointcloud_subsampling*.so with g++ and
pybind11.  The reference's build system (scikit-build + Conan + CMake, needing Eigen 3.4 / cgal / fmt
from the network) is not run; Eigen is replaced by the stand-in headers of oracle/eigen_standin/
(see the comment at the top of Eigen/Core there for what they assume).

Only possible in the build container (the GPU box has no /root/reference; the built .so travels).
Used by tests/ to pin oracle/native_oracle.c (count, order and bits of the output).
Called by __graft_entry__.build() when the reference is present."""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = Path(os.environ.get("MDI_REFERENCE_ROOT", "/root/reference")) / "native_modules"
SRC = REF / "subsampling" / "src"
OUT_DIR = HERE / "_ref"
OUT = OUT_DIR / ("_pointcloud_subsampling" + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))


def available() -> bool:
    return (SRC / "impl.cpp").exists()


def built() -> bool:
    return OUT.exists()


def build(force: bool = False) -> Path:
    if not available():
        if built():
            return OUT
        raise RuntimeError(f"reference native sources not present at {SRC}")
    srcs = [SRC / "impl.cpp", SRC / "pointcloud_subsampling.cpp"]
    deps = srcs + list(SRC.glob("*.h")) + list((HERE / "eigen_standin").rglob("*")) + [Path(__file__)]
    if not force and built() and all(OUT.stat().st_mtime >= d.stat().st_mtime for d in deps if d.is_file()):
        return OUT
    import pybind11
    OUT_DIR.mkdir(exist_ok=True)
    cmd = ["g++", "-std=c++20", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math", "-fvisibility=hidden",
           "-w",
           f"-I{HERE / 'eigen_standin'}",            # Eigen/Core, Eigen/Dense, pybind11/eigen.h stand-ins (first!)
           f"-I{REF / 'common' / 'include'}",       # <common/util.h>, <common/py_output_array.h>
           f"-I{SRC}",
           f"-I{pybind11.get_include()}", f"-I{sysconfig.get_paths()['include']}",
           *map(str, srcs), "-o", str(OUT)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building the reference native module failed:\n" + r.stderr[-6000:])
    return OUT


def load():
    """The compiled reference module (`_pointcloud_subsampling.subsample_pointcloud`)."""
    import importlib.util
    path = build()
    spec = importlib.util.spec_from_file_location("_pointcloud_subsampling", path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
