"""CLI (the reference's missing main.cpp) on the CPU path."""
import json
import subprocess
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
TINY = ["--num-char", "64", "--emb-dim", "64", "--hidden", "128", "--max-len", "6"]


def _run(*args):
    return subprocess.run([sys.executable, "-m", "gru_mpi_cuda_b200.cli", *args], capture_output=True, text=True,
                          cwd=ROOT, timeout=300)


def test_make_weights_then_generate_on_cpu(tmp_path):
    w, out = tmp_path / "w.bin", tmp_path / "names.bin"
    r = _run("--make-weights", str(w), *TINY)
    assert r.returncode == 0, r.stderr
    assert w.stat().st_size > 0
    r = _run("--params", str(w), "--n", "12", "--seed", "3", "--mode", "cpu", "--out", str(out), "--json", *TINY)
    assert r.returncode == 0, r.stderr
    rec = json.loads(r.stdout.strip().splitlines()[-1])
    assert rec["n_names"] == 12 and rec["names_per_sec"] > 0
    assert "Hello world, rank 0 out of 1" in r.stderr            # the reference's only log line
    data = torch.frombuffer(bytearray(out.read_bytes()), dtype=torch.uint8).view(12, 7)
    from gru_mpi_cuda_b200 import GRUConfig
    from gru_mpi_cuda_b200.models import oracle
    from gru_mpi_cuda_b200.utils import weights as W
    cfg = GRUConfig.tiny()
    want = oracle.generate(cfg, W.load(w, cfg), W.random_floats(12 * cfg.max_len, 3))
    assert torch.equal(data, want)


def test_cli_requires_weights():
    r = _run("--n", "4", "--mode", "cpu")
    assert r.returncode != 0
