"""The C ABI from a plain C program (examples/namegen_demo.c): the header must
be valid C, the library must link without CUDA headers, and the output must
match the Python-side oracle."""
import shutil
import subprocess
from pathlib import Path

import pytest
import torch

from gru_mpi_cuda_b200 import GRUConfig, _build, _native
from gru_mpi_cuda_b200.models import oracle
from gru_mpi_cuda_b200.utils import weights as W

ROOT = Path(__file__).resolve().parent.parent


def _compile(tmp_path):
    _native.lib()
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("no gcc")
    exe = tmp_path / "namegen_demo"
    lib_dir = _build.LIB_PATH.parent
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-Werror", "-O1", "-I", str(ROOT / "include"),
                        str(ROOT / "examples" / "namegen_demo.c"), "-L", str(lib_dir), "-lb200gru",
                        f"-Wl,-rpath,{lib_dir}", "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def _expected(n):
    cfg = GRUConfig.tiny()
    return oracle.generate(cfg, W.random_init(cfg, 11), W.random_floats(n * cfg.max_len, 5))


def test_c_demo_cpu_or_gpu_exact(tmp_path):
    """Runs anywhere: the demo uses the CPU oracle without a GPU and the exact
    kernels with one; both must reproduce the oracle bit for bit."""
    exe = _compile(tmp_path)
    out = tmp_path / "names.bin"
    r = subprocess.run([str(exe), "20", "exact", str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    got = torch.frombuffer(bytearray(out.read_bytes()), dtype=torch.uint8).view(20, 7)
    assert torch.equal(got, _expected(20))


@pytest.mark.gpu
def test_c_demo_tensor_core_path(tmp_path):
    exe = _compile(tmp_path)
    out = tmp_path / "names.bin"
    r = subprocess.run([str(exe), "300", "tc", str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "tcgen05" in r.stdout
    got = torch.frombuffer(bytearray(out.read_bytes()), dtype=torch.uint8).view(300, 7)
    assert (got == _expected(300)).all(dim=1).float().mean() > 0.8
