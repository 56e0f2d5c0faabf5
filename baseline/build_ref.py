"""Builds the reference arm: `python -m baseline.build_ref`.

`pip install /root/reference` is impossible (the reference is a C++/CUDA coursework tree with no
setup.py / pyproject.toml; recorded in DESIGN.md), and its own Makefile needs mpic++, Armadillo and
the CUDA samples headers, none of which exist on this image.  So the reference SOURCES are copied
verbatim into baseline/_ref/fpcode/ (git-ignored, shipped to the GPU box) and compiled unmodified
for sm_100 against three small stand-ins in baseline/shims/ (a mini Armadillo, an MPI subset over
shared memory, checkCudaErrors), plus baseline/harness.cpp which calls the reference's public
training API.  Flags follow fpcode/Makefile: -O2, one GPU architecture.
"""
from __future__ import annotations

import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_SRC = Path("/root/reference/fpcode")
DST = HERE / "_ref" / "fpcode"
LIB = HERE / "_ref" / "libref_mlp.so"
FILES = ["gpu_func.cu", "neural_network.cpp", "inc/gpu_func.h", "inc/neural_network.h",
         "utils/common.cpp", "utils/common.h"]


def copy_sources() -> bool:
    if not REF_SRC.exists():
        return all((DST / f).exists() for f in FILES)
    for f in FILES:
        (DST / f).parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(REF_SRC / f, DST / f)
    return True


def build(force: bool = False) -> Path | None:
    if LIB.exists() and not force:
        return LIB
    if not copy_sources():
        return None
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(nvcc).exists():
        return None
    cmd = [nvcc, "-O2", "-arch=sm_100", "-std=c++14", "-Xcompiler", "-fPIC", "-shared",
           "-Wno-deprecated-declarations", "-w",
           "-I", str(HERE / "shims"), "-I", str(DST / "inc"), "-I", str(DST),
           "-o", str(LIB),
           str(DST / "gpu_func.cu"), str(DST / "neural_network.cpp"), str(DST / "utils" / "common.cpp"),
           str(HERE / "harness.cpp"), str(HERE / "shims" / "mpi_shim.cpp"), "-lrt", "-lpthread"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        return None
    return LIB


if __name__ == "__main__":
    out = build(force="--force" in sys.argv)
    print("built" if out else "FAILED", out)
    raise SystemExit(0 if out else 1)
