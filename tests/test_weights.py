"""Weight-file layout / round trip (reference: namegensf.cu:370-407)."""
import torch

from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.utils import weights as W


def test_default_shape_byte_size():
    cfg = GRUConfig()
    assert W.param_count(cfg) == 11_415_808          # 45,663,232 bytes
    lay = W.layout(cfg)
    assert len(lay) == 27
    names = [t.name for t in lay]
    assert names[:7] == ["character_embedding", "W_ir0", "W_iz0", "W_in0", "W_ir1", "W_iz1", "W_in1"]
    assert names[7:13] == ["W_hr0", "W_hz0", "W_hn0", "W_hr1", "W_hz1", "W_hn1"]
    assert names[13:19] == ["b_ir0", "b_iz0", "b_in0", "b_ir1", "b_iz1", "b_in1"]
    assert names[19:25] == ["b_hr0", "b_hz0", "b_hn0", "b_hr1", "b_hz1", "b_hn1"]
    assert names[25:] == ["W_fc", "b_fc"]
    # contiguous, no gaps
    off = 0
    for t in lay:
        assert t.offset == off
        off += t.size
    assert lay[1].rows == 1024 and lay[1].cols == 512      # W_ir0 [H,E]
    assert lay[4].cols == 1024                             # W_ir1 [H,H]
    assert lay[25].rows == 256 and lay[25].cols == 1024    # W_fc [C,H]


def test_round_trip(tmp_path, tiny_cfg):
    p = W.random_init(tiny_cfg, seed=5)
    f = tmp_path / "w.bin"
    W.save(f, p)
    assert f.stat().st_size == p.numel() * 4
    q = W.load(f, tiny_cfg)
    assert torch.equal(p, q)
    parts = W.split(tiny_cfg, q)
    assert parts["W_hn1"].shape == (tiny_cfg.hidden, tiny_cfg.hidden)
    assert parts["b_fc"].shape == (tiny_cfg.num_char,)


def test_wrong_size_rejected(tmp_path, tiny_cfg):
    import pytest
    f = tmp_path / "w.bin"
    W.save(f, torch.zeros(10))
    with pytest.raises(ValueError):
        W.load(f, tiny_cfg)


def test_missing_file_is_an_error(tiny_cfg):
    import pytest
    from gru_mpi_cuda_b200._native import NativeError
    with pytest.raises(NativeError):
        W.load("/nonexistent/weights.bin", tiny_cfg)


def test_torch_state_round_trip(tiny_cfg):
    p = W.random_init(tiny_cfg, seed=6)
    sd = W.to_torch_state(tiny_cfg, p)
    assert sd["gru.weight_ih_l0"].shape == (3 * tiny_cfg.hidden, tiny_cfg.emb_dim)
    assert torch.equal(W.from_torch_state(tiny_cfg, sd), p)


def test_random_init_is_deterministic(tiny_cfg):
    assert torch.equal(W.random_init(tiny_cfg, 1), W.random_init(tiny_cfg, 1))
    assert not torch.equal(W.random_init(tiny_cfg, 1), W.random_init(tiny_cfg, 2))
    r = W.random_floats(1000, 3)
    assert r.min() >= 0 and r.max() < 1


def test_three_layer_layout():
    cfg = GRUConfig(num_char=32, emb_dim=16, hidden=64, num_layers=3)
    lay = W.layout(cfg)
    assert len(lay) == 3 + 12 * 3
    assert [t.name for t in lay][1:10] == ["W_ir0", "W_iz0", "W_in0", "W_ir1", "W_iz1", "W_in1",
                                           "W_ir2", "W_iz2", "W_in2"]
