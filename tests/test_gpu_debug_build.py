"""The checking build (B200GRU_DEBUG=1 -> libb200gru_debug.so) stands in for
compute-sanitizer, which this GPU pool refuses to run: device assertions on
computed offsets / ring and TMEM stage indices / sampled ids, guard bands around
every model buffer (verified after each generate and at destroy), and spin loops
that trap after 20 s instead of hanging.  This test runs the shape sweep of the
other GPU tests under it, in a child process so the release library of the test
session stays untouched."""
import os
import subprocess
import sys
import textwrap
from pathlib import Path

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(900)]
ROOT = Path(__file__).resolve().parent.parent

SCRIPT = textwrap.dedent("""
    import sys
    sys.path.insert(0, %r)
    import torch
    from gru_mpi_cuda_b200 import GRUConfig, _native
    from gru_mpi_cuda_b200.models import NameGenerator, oracle
    from gru_mpi_cuda_b200.parallel import partition
    from gru_mpi_cuda_b200.utils import weights as W

    assert _native.lib().b200gru_debug_build() == 1, "release library loaded"
    assert "libb200gru_debug.so" in open("/proc/self/maps").read()
    shapes = [
        GRUConfig.tiny(),
        GRUConfig(num_char=64, emb_dim=32, hidden=64, num_layers=1, max_len=5),
        GRUConfig(num_char=48, emb_dim=24, hidden=192, num_layers=3, max_len=4),
        GRUConfig(num_char=256, emb_dim=64, hidden=320, num_layers=2, max_len=3, sos=5, eos=9),
        GRUConfig(num_char=32, emb_dim=16, hidden=128, num_layers=4, max_len=3),
        GRUConfig(num_char=27, emb_dim=5, hidden=32, num_layers=2, max_len=4),      # exact path only
        GRUConfig(max_len=3),                                                       # default widths
    ]
    for cfg in shapes:
        p = W.random_init(cfg, 40 + cfg.hidden)
        for n in (1, 129, 700):
            r = W.random_floats(n * cfg.max_len, 41)
            want = oracle.generate(cfg, p, r) if cfg.hidden <= 320 else None
            with NameGenerator(cfg, p, device=0, max_batch=n) as g:
                ex = g.generate(n, r, mode="exact")
                if want is not None:
                    assert torch.equal(ex, want), (cfg, n)
                if cfg.hidden %% 64 == 0 and cfg.num_char %% 16 == 0:
                    g.set_path("multi")
                    a = g.generate(n, r)
                    assert (a[:, -1] == 0).all()
                    if g.set_path("auto"):
                        g.set_path("persistent")
                        b = g.generate(n, r)
                        assert (b[:, 0] == ex[:, 0]).float().mean() >= (0.9 if n > 100 else 0.0)
                g.check()
    # logical ranks with root-buffer offsets (the DP addressing) under the checks
    cfg = GRUConfig.tiny(); p = W.random_init(cfg, 11); n, T = 203, cfg.max_len
    rf = W.random_floats(n * T, 31)
    with NameGenerator(cfg, p, device=0, max_batch=n) as g:
        for path in ("multi", "persistent"):
            g.set_path(path)
            single = g.generate(n, rf)
            rf_dev = rf.cuda(); out_dev = torch.zeros(n, T + 1, dtype=torch.uint8, device="cuda")
            for rank in range(4):
                sl = partition(n, 4, rank)
                g.generate_device(rf_dev[sl.start * T:], sl.count, out_dev=out_dev[sl.start:])
            torch.cuda.synchronize()
            assert torch.equal(out_dev.cpu()[:200], single[:200]), path
        g.check()
    print("DEBUG_BUILD_SWEEP_OK")
""") % str(ROOT)


def test_shape_sweep_under_the_checking_build():
    env = dict(os.environ, B200GRU_DEBUG="1")
    res = subprocess.run([sys.executable, "-c", SCRIPT], capture_output=True, text=True, timeout=850, env=env, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "DEBUG_BUILD_SWEEP_OK" in res.stdout
    assert "debug assert" not in res.stdout + res.stderr
