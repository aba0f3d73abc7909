"""The persistent weight-stationary kernel (csrc/gru_persist_sm100.cu): one
launch per generate() for small batches.  Checked against the CPU oracle, the
fp32 torch model and the multi-launch tensor-core path; every test has a
timeout because a cooperative kernel whose CTAs wait on each other must fail
loudly, not hang (the kernel's own watchdog traps after ~10 s)."""
import pytest
import torch

from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.models import NameGenerator, TorchNameGen, oracle
from gru_mpi_cuda_b200.parallel import partition
from gru_mpi_cuda_b200.utils import weights as W

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


@pytest.fixture(scope="module")
def gen_tiny(tiny_cfg, tiny_params):
    g = NameGenerator(tiny_cfg, tiny_params, device=0, max_batch=1024, mode="tc")
    assert g.set_path("persistent"), "persistent kernel must be available for the tiny shape"
    yield g
    g.finalize()


@pytest.mark.parametrize("n", [1, 128, 129, 257, 700, 1024])
def test_persistent_close_to_oracle(gen_tiny, tiny_cfg, tiny_params, n):
    r = W.random_floats(n * tiny_cfg.max_len, 50 + n)
    want = oracle.generate(tiny_cfg, tiny_params, r)
    got = gen_tiny.generate(n, r)
    assert gen_tiny.last_launches == 1          # the whole call is one kernel
    again = gen_tiny.generate(n, r)
    assert torch.equal(got, again)
    assert got.shape == want.shape
    assert (got[:, -1] == 0).all()
    if n >= 128:
        assert (got[:, 0] == want[:, 0]).float().mean() > 0.98
        assert (got == want).all(dim=1).float().mean() > 0.80


def test_persistent_agrees_with_multi_launch_path(gen_tiny, tiny_cfg):
    n = 1000
    r = W.random_floats(n * tiny_cfg.max_len, 61)
    pers = gen_tiny.generate(n, r)
    gen_tiny.set_path("multi")
    multi = gen_tiny.generate(n, r)
    assert gen_tiny.last_launches > 1
    gen_tiny.set_path("persistent")
    # same bf16 operands and fp32 accumulation; only the place where the layer-0
    # table row is added differs, so almost every name is the same
    assert (pers == multi).all(dim=1).float().mean() > 0.93


def test_persistent_logits_against_fp32_reference(gen_tiny, tiny_cfg, tiny_params):
    n = 256
    r = W.random_floats(n * tiny_cfg.max_len, 62)
    gen_tiny.enable_logits_tap(True)
    got = gen_tiny.generate(n, r)
    logits = gen_tiny.debug_logits(n).cpu()
    gen_tiny.enable_logits_tap(False)
    model = TorchNameGen.from_flat(tiny_cfg, tiny_params)
    chars = torch.full((n,), tiny_cfg.sos, dtype=torch.long)
    h = torch.zeros(tiny_cfg.num_layers, n, tiny_cfg.hidden)
    alive = torch.ones(n, dtype=torch.bool)
    for t in range(tiny_cfg.max_len):
        ref, h = model.step(chars, h)
        chars = got[:, t].long()
        if t < tiny_cfg.max_len - 1:
            alive &= chars != tiny_cfg.eos
    assert alive.sum() > n // 2
    err = (logits[alive] - ref[alive]).abs().max().item()
    assert err < 0.05, err


def test_persistent_eos_at_step_zero(tiny_cfg):
    p = W.random_init(tiny_cfg, 0)
    W.split(tiny_cfg, p)["b_fc"][tiny_cfg.eos] = 30.0
    r = torch.full((300 * tiny_cfg.max_len,), 0.5)
    with NameGenerator(tiny_cfg, p, device=0, max_batch=512) as g:
        g.set_path("persistent")
        out = g.generate(300, r)
        assert (out[:, 0] == tiny_cfg.eos).all()
        assert (out[:, 1:] == 0).all()


def test_persistent_mid_sequence_eos_matches_exact(tiny_cfg):
    """EOS made likely (not certain): names end at different steps, whole
    128-name tiles die early and are skipped; rows must stay NUL after EOS and
    follow the exact path closely."""
    p = W.random_init(tiny_cfg, 3)
    W.split(tiny_cfg, p)["b_fc"][tiny_cfg.eos] = 3.5
    n = 640
    r = W.random_floats(n * tiny_cfg.max_len, 63)
    with NameGenerator(tiny_cfg, p, device=0, max_batch=n) as g:
        g.set_path("persistent")
        out = g.generate(n, r)
        ex = g.generate(n, r, mode="exact")
    T = tiny_cfg.max_len
    ended = 0
    for row in out[:, :T].tolist():
        if tiny_cfg.eos in row:
            k = row.index(tiny_cfg.eos)
            ended += 1
            assert all(v == 0 for v in row[k + 1:]), row
    assert ended > n // 2
    assert (out == ex).all(dim=1).float().mean() > 0.85


def test_persistent_default_shape_vs_exact():
    cfg = GRUConfig()
    p = W.random_init(cfg, 7)
    n = 384
    r = W.random_floats(n * cfg.max_len, 8)
    with NameGenerator(cfg, p, device=0, max_batch=n, mode="tc") as g:
        assert g.set_path("persistent")
        tc = g.generate(n, r)
        assert g.last_launches == 1
        ex = g.generate(n, r, mode="exact")
        g.set_path("multi")
        ml = g.generate(n, r)
    assert (tc[:, 0] == ex[:, 0]).float().mean() > 0.99
    assert (tc == ex).all(dim=1).float().mean() > 0.95
    assert (tc == ml).all(dim=1).float().mean() > 0.95


@pytest.mark.parametrize("world", [2, 8])
def test_persistent_logical_ranks_bit_identical(gen_tiny, tiny_cfg, world):
    """DP semantics: each logical rank's slice, run as its own call with
    root-buffer offsets, reproduces the rows of the single-rank run exactly."""
    cfg = tiny_cfg
    n, T = 203, cfg.max_len
    rf = W.random_floats(n * T, 64)
    single = gen_tiny.generate(n, rf)
    rf_dev = rf.cuda()
    out_dev = torch.zeros(n, T + 1, dtype=torch.uint8, device="cuda")
    for rank in range(world):
        sl = partition(n, world, rank)
        gen_tiny.generate_device(rf_dev[sl.start * T:], sl.count, out_dev=out_dev[sl.start:])
    torch.cuda.synchronize()
    covered = (n // world) * world
    assert torch.equal(out_dev.cpu()[:covered], single[:covered])


def test_auto_path_switches_on_batch_size(tiny_cfg, tiny_params):
    with NameGenerator(tiny_cfg, tiny_params, device=0, max_batch=2048) as g:
        assert g.set_path("auto", persist_max_names=256)
        g.set_compaction(False)        # its extra launches are counted in last_launches
        r = W.random_floats(2048 * tiny_cfg.max_len, 65)
        g.generate(200, r)
        assert g.last_launches == 1
        g.generate(2000, r)
        assert g.last_launches == 1 + tiny_cfg.max_len * (tiny_cfg.num_layers + 1)


def test_persistent_unavailable_shape_is_reported():
    """Three layers at H = 1024 do not fit in shared memory: auto falls back to
    the multi-launch path, forcing the persistent path is an error."""
    from gru_mpi_cuda_b200._native import NativeError
    cfg = GRUConfig(num_char=64, emb_dim=32, hidden=1024, num_layers=3, max_len=3)
    p = W.random_init(cfg, 1)
    r = W.random_floats(64 * cfg.max_len, 2)
    with NameGenerator(cfg, p, device=0, max_batch=64) as g:
        assert not g.set_path("auto")
        g.generate(64, r)
        assert g.last_launches > 1
        g.set_path("persistent")
        with pytest.raises(NativeError, match="persistent kernel cannot run"):
            g.generate(64, r)
