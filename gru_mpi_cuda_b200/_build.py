"""In-tree build of libb200gru.so (nvcc, sm_100a only).

No torch cpp_extension: the library has a plain C ABI (include/b200gru.h) and
is loaded with ctypes, so a compile is a few seconds per file and the built
`.so` travels with the source tree to the GPU box.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
LIB_DIR = Path(__file__).resolve().parent / "lib"
OBJ_DIR = ROOT / "build" / "obj"
# B200GRU_DEBUG=1 selects the checking build (device assertions, guard bands,
# bounded spin loops; see include/b200gru.h): a second library next to the
# release one, so both can travel to the GPU box together.
DEBUG = os.environ.get("B200GRU_DEBUG") == "1"
LIB_PATH = LIB_DIR / ("libb200gru_debug.so" if DEBUG else "libb200gru.so")

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-ffp-contract=off,-Wall",
          "-I", str(ROOT / "include")] + (["-DB200GRU_DEBUG"] if DEBUG else [])
# fmad stays ON globally (tensor-core epilogues want it); the exact kernel uses
# explicit __fmaf_rn/__fmul_rn/__fadd_rn so contraction cannot change it.
CU_FLAGS = ["--expt-relaxed-constexpr", "-Xptxas", "-v"]


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(exe).exists():
        raise RuntimeError("nvcc not found; libb200gru.so cannot be built")
    return exe


def sources() -> list[Path]:
    return sorted(list(CSRC.glob("*.cpp")) + list(CSRC.glob("*.cu")))


def _headers_digest() -> str:
    h = hashlib.sha1()
    for p in sorted(list(CSRC.glob("*.h")) + list(CSRC.glob("*.hpp")) + list(CSRC.glob("*.cuh"))
                    + list((ROOT / "include").glob("*.h"))):
        h.update(p.read_bytes())
    h.update(" ".join(ARCH_FLAGS + COMMON + CU_FLAGS).encode())
    return h.hexdigest()[:16]


def _compile_one(src: Path, digest: str, verbose: bool) -> Path:
    obj = OBJ_DIR / (src.name + "." + digest + ".o")
    if obj.exists() and obj.stat().st_mtime >= src.stat().st_mtime:
        return obj
    cmd = [nvcc(), *ARCH_FLAGS, *COMMON]
    if src.suffix == ".cu":
        cmd += CU_FLAGS
    cmd += ["-c", str(src), "-o", str(obj)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = OBJ_DIR / (src.name + ".log")
    log.write_text(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src.name}:\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(f"[build] {src.name} ok", file=sys.stderr)
    return obj


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every source under csrc/ for sm_100a and link libb200gru.so.

    Serialised across processes with a file lock: under torchrun every rank
    imports the package at the same moment."""
    import fcntl
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    with open(OBJ_DIR / ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not needs_build():
                return LIB_PATH          # another process built it while we waited
            return _build_locked(force, verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(force: bool, verbose: bool) -> Path:
    digest = _headers_digest()
    srcs = sources()
    if force:
        for o in OBJ_DIR.glob("*.o"):
            o.unlink()
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        objs = list(ex.map(lambda s: _compile_one(s, digest, verbose), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if force or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < newest:
        tmp = LIB_PATH.with_suffix(".so.tmp%d" % os.getpid())
        cmd = [nvcc(), *ARCH_FLAGS, "-shared", "-o", str(tmp), *map(str, objs),
               "-cudart", "static", "-lpthread", "-ldl", "-lrt"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
        os.replace(tmp, LIB_PATH)
        if verbose:
            print(f"[build] linked {LIB_PATH}", file=sys.stderr)
    return LIB_PATH


def needs_build() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    deps = sources() + list(CSRC.glob("*.h")) + list(CSRC.glob("*.hpp")) + \
        list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    return any(p.stat().st_mtime > t for p in deps)


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    print(LIB_PATH)
