"""Multi-rank behaviour (SURVEY §4 'Distributed' tier).

CPU: 2- and 3-rank gloo runs of the reference's scatter/generate/gather flow.
GPU: logical ranks emulated on one GPU (slices must be bit-identical to the
1-rank run) and, when >= 2 GPUs are visible, real ranks with in-kernel P2P.
"""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
WORKER = str(ROOT / "tests" / "dist_worker.py")


def _torchrun(nproc, *args, timeout=300):
    port = 29500 + (os.getpid() * 7 + nproc) % 2000
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), WORKER, *map(str, args)]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


@pytest.mark.parametrize("world,n", [(2, 10), (3, 10)])
def test_cpu_scatter_gather_semantics(world, n):
    r = _torchrun(world, "cpu", n)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert f"CPU_DP_OK world={world}" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_logical_ranks_on_one_gpu_bit_identical(world, tiny_cfg, tiny_params):
    """Each logical rank's slice, run separately with root-buffer offsets,
    reproduces the corresponding rows of the single-rank run exactly."""
    from gru_mpi_cuda_b200.models import NameGenerator
    from gru_mpi_cuda_b200.parallel import partition
    from gru_mpi_cuda_b200.utils import weights as W
    cfg = tiny_cfg
    n = 203
    T = cfg.max_len
    rf = W.random_floats(n * T, 31)
    with NameGenerator(cfg, tiny_params, device=0, max_batch=n) as g:
        for mode in ("exact", "tc"):
            single = g.generate(n, rf, mode=mode)
            rf_dev = rf.cuda()
            out_dev = torch.zeros(n, T + 1, dtype=torch.uint8, device="cuda")
            for rank in range(world):
                sl = partition(n, world, rank)
                g.generate_device(rf_dev[sl.start * T:], sl.count, out_dev=out_dev[sl.start:], mode=mode)
            torch.cuda.synchronize()
            covered = (n // world) * world
            assert torch.equal(out_dev.cpu()[:covered], single[:covered]), mode
            assert (out_dev.cpu()[covered:] == 0).all()


@pytest.mark.gpu
@pytest.mark.parametrize("method,n", [("p2p", 1001), ("nccl", 1001), ("p2p", 1)])
def test_real_ranks_p2p(method, n):
    """n = 1 with >= 2 ranks: JPP = 0, every rank is idle and the row stays zero."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(torch.cuda.device_count(), 8)
    r = _torchrun(world, "gpu", n, method, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("GPU_DP_OK") == 3


@pytest.mark.gpu
@pytest.mark.parametrize("hidden,n", [(256, 600), (512, 130)])
def test_tensor_parallel_matches_single_gpu(hidden, n):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if hidden == 256 else min(4, torch.cuda.device_count())
    world = max(w for w in (1, 2, 4, 8) if w <= world and hidden % (64 * w) == 0)
    r = _torchrun(world, "tp", n, hidden, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "GPU_TP_OK" in r.stdout


@pytest.mark.gpu
def test_tensor_parallel_h8192_on_8_gpus():
    """BASELINE.json config 4 at its real shape: hidden 8192 sharded over 8 GPUs
    (1024 units per rank), small batch, against the single-GPU tensor-core path
    on the same weights.  Batch sizes vary between calls and every call has its
    own random floats (see dist_worker.run_tp)."""
    if torch.cuda.device_count() < 8:
        pytest.skip("needs 8 GPUs")
    r = _torchrun(8, "tp", 300, 8192, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "GPU_TP_OK world=8 hidden=8192" in r.stdout
