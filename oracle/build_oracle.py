"""TEST INFRASTRUCTURE -- builds oracle/native_oracle.c (the C restatement of the reference's
native module) into oracle/_build/libmdi_native_oracle.so with gcc.  Called by
__graft_entry__.build(); loading the result is reserved to tests/, smoke() and bench.py's CPU legs.
The reference module itself is unbuildable here (Eigen + Conan + CMake), so there is no oracle/_ref."""
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = HERE / "native_oracle.c"
OUT = HERE / "_build" / "libmdi_native_oracle.so"


def build() -> Path:
    OUT.parent.mkdir(exist_ok=True)
    if OUT.exists() and OUT.stat().st_mtime >= SRC.stat().st_mtime:
        return OUT
    subprocess.run(["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math",
                    str(SRC), "-o", str(OUT)], check=True)
    return OUT


if __name__ == "__main__":
    print(build())
