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


def test_torch_state_dict_round_trip_through_the_cli(tmp_path):
    from gru_mpi_cuda_b200 import GRUConfig
    from gru_mpi_cuda_b200.models import TorchNameGen
    from gru_mpi_cuda_b200.utils import weights as W
    cfg = GRUConfig.tiny()
    model = TorchNameGen(cfg)                       # stock torch init
    pt, w, pt2 = tmp_path / "m.pt", tmp_path / "w.bin", tmp_path / "m2.pt"
    torch.save(model.state_dict(), pt)
    assert _run("--import-torch", str(pt), str(w), *TINY).returncode == 0
    assert w.stat().st_size == W.param_count(cfg) * 4
    assert _run("--export-torch", str(w), str(pt2), *TINY).returncode == 0
    sd = torch.load(pt2, weights_only=True)
    for k, v in model.state_dict().items():
        assert torch.equal(sd[k], v), k
