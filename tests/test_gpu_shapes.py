"""Shape coverage on the GPU: layer counts, non-power-of-two hidden sizes,
alphabet sizes that are not a multiple of the 64-wide one-hot block, odd batch
sizes — exact path bit-identical to the oracle, tensor-core path close to it."""
import pytest
import torch

from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200._native import NativeError
from gru_mpi_cuda_b200.models import NameGenerator, oracle
from gru_mpi_cuda_b200.utils import weights as W

pytestmark = pytest.mark.gpu

SHAPES = [
    GRUConfig(num_char=64, emb_dim=32, hidden=64, num_layers=1, max_len=5),
    GRUConfig(num_char=48, emb_dim=24, hidden=192, num_layers=3, max_len=4),     # C -> kx 64, H not 2^k
    GRUConfig(num_char=256, emb_dim=64, hidden=320, num_layers=2, max_len=3, sos=5, eos=9),
    GRUConfig(num_char=32, emb_dim=16, hidden=128, num_layers=4, max_len=3),
]


@pytest.mark.parametrize("cfg", SHAPES, ids=lambda c: f"C{c.num_char}E{c.emb_dim}H{c.hidden}L{c.num_layers}")
@pytest.mark.parametrize("n", [1, 129, 700])
def test_shapes_exact_and_tc(cfg, n):
    p = W.random_init(cfg, 40 + cfg.hidden)
    r = W.random_floats(n * cfg.max_len, 41)
    want = oracle.generate(cfg, p, r)
    with NameGenerator(cfg, p, device=0, max_batch=n) as g:
        ex = g.generate(n, r, mode="exact")
        tc = g.generate(n, r, mode="tc")
        tc2 = g.generate(n, r, mode="tc")          # second call replays the CUDA graph
    assert torch.equal(ex, want)
    assert torch.equal(tc, tc2)
    assert (tc[:, 0] == want[:, 0]).float().mean() >= 0.95
    assert (tc == want).all(dim=1).float().mean() >= (0.8 if n > 1 else 0.0)
    assert (tc[:, -1] == 0).all()


def test_exact_path_with_unaligned_weight_offsets():
    """num_char*emb_dim = 135 floats puts every later tensor of the weight file on
    an address that is not 16-byte aligned while K % 4 == 0 for hidden = 32: the
    SGEMM must then take its scalar loads, not fault on a float4."""
    cfg = GRUConfig(num_char=27, emb_dim=5, hidden=32, num_layers=2, max_len=4, sos=0, eos=1)
    p = W.random_init(cfg, 3)
    for n in (1, 77):
        r = W.random_floats(n * cfg.max_len, 4)
        with NameGenerator(cfg, p, device=0, max_batch=n, mode="exact") as g:
            assert torch.equal(g.generate(n, r), oracle.generate(cfg, p, r))


def test_unsupported_tc_shape_falls_back_loudly():
    """hidden not a multiple of 64: the tensor-core path refuses, the exact path runs."""
    cfg = GRUConfig(num_char=32, emb_dim=16, hidden=40, num_layers=2, max_len=3)
    p = W.random_init(cfg, 1)
    r = W.random_floats(10 * cfg.max_len, 2)
    with NameGenerator(cfg, p, device=0, max_batch=16, mode="exact") as g:
        assert torch.equal(g.generate(10, r), oracle.generate(cfg, p, r))
        with pytest.raises(NativeError, match="tensor-core mode needs"):
            g.generate(10, r, mode="tc")


def test_too_many_names_is_an_error(tiny_cfg, tiny_params):
    with NameGenerator(tiny_cfg, tiny_params, device=0, max_batch=8) as g:
        with pytest.raises(ValueError):
            g.generate_device(torch.zeros(1000, device="cuda"), 9)


def test_zero_names_is_a_no_op(tiny_cfg, tiny_params):
    with NameGenerator(tiny_cfg, tiny_params, device=0, max_batch=8) as g:
        for mode in ("exact", "tc"):
            out = g.generate(0, torch.zeros(0), mode=mode)
            assert out.shape == (0, tiny_cfg.max_len + 1)
