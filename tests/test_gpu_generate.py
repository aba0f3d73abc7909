"""GPU kernels vs the CPU oracles (SURVEY §4 'Kernel' tier)."""
import os

import pytest
import torch

from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.models import NameGenerator, TorchNameGen, oracle
from gru_mpi_cuda_b200.utils import weights as W

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gen_tiny(tiny_cfg, tiny_params):
    g = NameGenerator(tiny_cfg, tiny_params, device=0, max_batch=1024, mode="tc")
    yield g
    g.finalize()


def test_native_library_is_loaded():
    from gru_mpi_cuda_b200 import _native
    assert _native.lib().b200gru_version().decode().startswith("b200gru")
    maps = open("/proc/self/maps").read()
    assert "libb200gru.so" in maps


@pytest.mark.parametrize("m,n,k", [(128, 64, 64), (128, 192, 128), (300, 256, 256), (256, 512, 1024), (1, 16, 64)])
def test_tcgen05_gemm_probe(m, n, k):
    from gru_mpi_cuda_b200.ops import gemm_bf16_tn
    torch.manual_seed(m + n + k)
    a = (torch.randn(m, k, device="cuda") * 0.5).bfloat16()
    b = (torch.randn(n, k, device="cuda") * 0.5).bfloat16()
    d = gemm_bf16_tn(a, b)
    ref = a.float() @ b.float().t()
    err = (d - ref).abs().max().item()
    assert err < 1e-2 * max(1.0, ref.abs().max().item()), err


def test_exact_matches_oracle_bitwise(gen_tiny, tiny_cfg, tiny_params):
    n = 300
    r = W.random_floats(n * tiny_cfg.max_len, 21)
    want, wl = oracle.generate(tiny_cfg, tiny_params, r, return_logits=True)
    got = gen_tiny.generate(n, r, mode="exact")
    assert torch.equal(got, want)
    # last-step logits are bit-identical too
    gl = gen_tiny.debug_logits(n).cpu()
    assert torch.equal(gl, wl[:, -1])


def test_exact_repeatable_and_chunked(tiny_cfg, tiny_params):
    """n > exact chunk is not needed at tiny sizes; instead check two calls and a
    smaller max_batch give identical rows."""
    n = 64
    r = W.random_floats(n * tiny_cfg.max_len, 22)
    with NameGenerator(tiny_cfg, tiny_params, device=0, max_batch=64, mode="exact") as g:
        a = g.generate(n, r)
        b = g.generate(n, r)
    assert torch.equal(a, b)
    assert torch.equal(a, oracle.generate(tiny_cfg, tiny_params, r))


def test_tc_close_to_oracle(gen_tiny, tiny_cfg, tiny_params):
    n = 1000
    r = W.random_floats(n * tiny_cfg.max_len, 23)
    want, wl = oracle.generate(tiny_cfg, tiny_params, r, return_logits=True)
    got = gen_tiny.generate(n, r, mode="tc")
    assert got.shape == want.shape
    # first character depends only on step-0 math: near-perfect agreement
    first = (got[:, 0] == want[:, 0]).float().mean().item()
    assert first > 0.98, first
    names = (got == want).all(dim=1).float().mean().item()
    assert names > 0.80, names
    assert (got[:, -1] == 0).all()


def test_tc_logits_against_fp32_reference(gen_tiny, tiny_cfg, tiny_params):
    """Teacher-forced check: run the torch fp32 model along the characters the
    tensor-core path actually sampled and compare final-step logits."""
    n = 256
    r = W.random_floats(n * tiny_cfg.max_len, 24)
    gen_tiny.enable_logits_tap(True)
    got = gen_tiny.generate(n, r, mode="tc")
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
    sel = alive
    assert sel.sum() > n // 2
    err = (logits[sel] - ref[sel]).abs().max().item()
    assert err < 0.05, err


def test_default_shape_tc_vs_exact():
    cfg = GRUConfig()
    p = W.random_init(cfg, 7)
    n = 512
    r = W.random_floats(n * cfg.max_len, 8)
    with NameGenerator(cfg, p, device=0, max_batch=n, mode="tc") as g:
        tc = g.generate(n, r)
        ex = g.generate(n, r, mode="exact")
    want = oracle.generate(cfg, p, r, n_names=16)
    assert torch.equal(ex[:16], want)
    # bounds derived from the measured logit error, see tests/test_gpu_accuracy.py
    assert (tc[:, 0] == ex[:, 0]).float().mean() > 0.99
    assert (tc == ex).all(dim=1).float().mean() > 0.95


def test_eos_stops_and_pads(tiny_cfg):
    p = W.random_init(tiny_cfg, 0)
    W.split(tiny_cfg, p)["b_fc"][tiny_cfg.eos] = 30.0
    r = torch.full((130 * tiny_cfg.max_len,), 0.5)
    with NameGenerator(tiny_cfg, p, device=0, max_batch=256) as g:
        for mode in ("exact", "tc"):
            out = g.generate(130, r, mode=mode)
            assert (out[:, 0] == tiny_cfg.eos).all(), mode
            assert (out[:, 1:] == 0).all(), mode


def test_trace_reports_every_launch(gen_tiny, tiny_cfg):
    n = 300
    r = W.random_floats(n * tiny_cfg.max_len, 25)
    gen_tiny.set_path("multi")      # the per-layer launches; the persistent path is one span
    gen_tiny.set_compaction(False)  # (its three launches per point share one span)
    gen_tiny.set_trace(True)
    gen_tiny.generate(n, r, mode="tc")
    tr = gen_tiny.trace()
    gen_tiny.set_trace(False)
    gen_tiny.set_compaction(True)
    gen_tiny.set_path("auto")
    assert len(tr) == gen_tiny.last_launches == 1 + tiny_cfg.max_len * (tiny_cfg.num_layers + 1)
    assert [k for k, _ in tr[:4]] == ["init", "layer0", "layerN", "fc_sample"]
    assert all(t >= 0 for _, t in tr)


def test_tc_character_distribution_matches_exact(gen_tiny, tiny_cfg):
    """Sampling amplifies 1-ulp CDF differences, so bf16 names are compared
    statistically: two-sample chi-square on the character histograms of the
    tensor-core and the exact path (same random floats)."""
    n = 1024
    r = W.random_floats(n * tiny_cfg.max_len, 26)
    tc = gen_tiny.generate(n, r, mode="tc")[:, : tiny_cfg.max_len].reshape(-1).long()
    ex = gen_tiny.generate(n, r, mode="exact")[:, : tiny_cfg.max_len].reshape(-1).long()
    C = tiny_cfg.num_char
    a = torch.bincount(tc, minlength=C).double()
    b = torch.bincount(ex, minlength=C).double()
    keep = (a + b) > 0
    chi2 = (((a - b) ** 2) / (a + b))[keep].sum().item()
    dof = int(keep.sum().item()) - 1
    # identical streams of random floats make the two samples strongly dependent,
    # so chi2 is far below dof; dof + 4*sqrt(2*dof) would already be generous
    assert chi2 < dof + 4 * (2 * dof) ** 0.5, (chi2, dof)
    assert (tc == ex).double().mean() > 0.9


@pytest.mark.parametrize("shape", ["tiny", "default"])
def test_compaction_of_finished_names_changes_nothing_but_the_time(shape, tiny_cfg):
    """EOS made likely: most names end early, the multi-launch path packs the
    survivors to the front after steps 1, 3, 5, ... and later kernels run fewer
    tiles.  Rows must be bit-identical to a run with compaction off (a row's
    arithmetic does not depend on where in the batch it sits)."""
    cfg = tiny_cfg if shape == "tiny" else GRUConfig()
    p = W.random_init(cfg, 3)
    W.split(cfg, p)["b_fc"][cfg.eos] = 3.0 if shape == "tiny" else 4.4     # P(EOS) ~ 0.25 per step
    n = 3000 if shape == "tiny" else 4096
    r = W.random_floats(n * cfg.max_len, 71)
    os.environ["B200GRU_COMPACT_MIN"] = "0"      # small calls skip compaction by default
    try:
        g = NameGenerator(cfg, p, device=0, max_batch=n)
    finally:
        del os.environ["B200GRU_COMPACT_MIN"]
    with g:
        g.set_path("multi")
        on = g.generate(n, r)
        launches_on = g.last_launches
        again = g.generate(n, r)           # graph replay, state fully reset per call
        g.set_compaction(False)
        off = g.generate(n, r)
        launches_off = g.last_launches
        ex = g.generate(n, r, mode="exact") if shape == "tiny" else None
    assert torch.equal(on, off)
    assert torch.equal(on, again)
    assert launches_on > launches_off
    T = cfg.max_len
    ended = (on[:, :T] == cfg.eos).any(dim=1).float().mean().item()
    assert ended > 0.5, ended
    if ex is not None:
        assert (on == ex).all(dim=1).float().mean() > 0.9
