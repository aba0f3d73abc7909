"""Build the UNMODIFIED reference (namegensf.cu) against the shim headers.

The reference ships no build system, no main.cpp, no namegen.h/util.h, and this
image has no MPI.  `pip install /root/reference` fails ("neither
setup.py nor pyproject.toml"), so the same-box baseline is this harness:
namegensf.cu is copied byte-for-byte into baseline/_ref/ and compiled for
sm_100a with baseline/shim/ on the include path.  The resulting binary
baseline/_ref/namegen_ref travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import hashlib
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_SRC = Path("/root/reference/namegensf.cu")
OUT = HERE / "_ref"
SHIM = HERE / "shim"
BIN = OUT / "namegen_ref"


def build(verbose: bool = True) -> Path | None:
    OUT.mkdir(exist_ok=True)
    dst = OUT / "namegensf.cu"
    if REF_SRC.exists():
        shutil.copyfile(REF_SRC, dst)          # unmodified copy
    if not dst.exists():
        return BIN if BIN.exists() else None
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    arch = ["-gencode", "arch=compute_100a,code=sm_100a"]
    cmds = [
        [nvcc, *arch, "-O3", "-w", "-I", str(SHIM), "-c", str(dst), "-o", str(OUT / "namegensf.o")],
        [nvcc, *arch, "-O3", "-I", str(SHIM), "-c", str(SHIM / "main.cpp"), "-o", str(OUT / "main.o")],
        [nvcc, "-O3", "-I", str(SHIM), "-c", str(SHIM / "mpi_shim.cpp"), "-o", str(OUT / "mpi_shim.o")],
        [nvcc, *arch, "-o", str(BIN), str(OUT / "namegensf.o"), str(OUT / "main.o"), str(OUT / "mpi_shim.o"),
         "-cudart", "static", "-lrt", "-lpthread"],
    ]
    for c in cmds:
        r = subprocess.run(c, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("reference build failed: " + " ".join(c) + "\n" + r.stdout + r.stderr)
    (OUT / "SHA256").write_text(hashlib.sha256(dst.read_bytes()).hexdigest() + "  namegensf.cu\n")
    if verbose:
        print(f"[baseline] built {BIN}", file=sys.stderr)
    return BIN


if __name__ == "__main__":
    print(build())
