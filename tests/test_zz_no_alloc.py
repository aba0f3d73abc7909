"""b200gru_gru_cell_fp32_noalloc, called through the C ABI.

This entry point exists for callers that cannot allocate GPU memory when they
call into the library — a host whose HBM is already handed out to another
pool, or a loop that cannot afford the device-wide sync of cudaMalloc/cudaFree.
Its GEMM scratch is a fixed 4 MiB array inside the library, shared by every
call on a GPU, so the things worth checking are: same result as the
allocating entry point, still the same on a second call that finds the first
call's leftovers in the scratch, and a clean refusal above the size limit.

The file name sorts last so that a `-x` run has finished every other test
before it gets here.
"""
import ctypes
from ctypes import c_int, c_void_p

import pytest
import torch

from gru_mpi_cuda_b200 import _build, _native

SCRATCH_BYTES = 4 << 20
ARGS = [c_void_p] * 7 + [c_int, c_int, c_int, c_void_p]


def _lib():
    _native.lib()                               # builds the library if the tree is newer
    lib = ctypes.CDLL(str(_build.LIB_PATH))     # then talk to it through the plain C ABI
    for name in ("b200gru_gru_cell_fp32", "b200gru_gru_cell_fp32_noalloc"):
        fn = getattr(lib, name)
        fn.restype = c_int
        fn.argtypes = ARGS
    lib.b200gru_last_error.restype = ctypes.c_char_p
    return lib


def _case(B, E, H, seed):
    g = torch.Generator().manual_seed(seed)
    mk = lambda *s, sc=1.0: (torch.randn(*s, generator=g) * sc).cuda()
    return dict(x=mk(B, E), h=mk(B, H, sc=0.5), w_ih=mk(3 * H, E, sc=0.2), w_hh=mk(3 * H, H, sc=0.2),
                b_ih=mk(3 * H, sc=0.1), b_hh=mk(3 * H, sc=0.1))


def _call(fn, c, B, E, H):
    out = torch.empty(B, H, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = fn(c["x"].data_ptr(), c["h"].data_ptr(), c["w_ih"].data_ptr(), c["w_hh"].data_ptr(),
            c["b_ih"].data_ptr(), c["b_hh"].data_ptr(), out.data_ptr(), B, E, H, c_void_p(stream))
    torch.cuda.synchronize()
    return rc, out


@pytest.mark.gpu
def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    lib = _lib()
    # two different shapes back to back: the second call starts from scratch
    # that still holds the first call's GEMM outputs
    for B, E, H, seed in [(96, 64, 128, 1), (33, 48, 256, 2)]:
        assert 2 * B * 3 * H * 4 <= SCRATCH_BYTES
        c = _case(B, E, H, seed)
        rc_a, want = _call(lib.b200gru_gru_cell_fp32, c, B, E, H)
        rc_b, got = _call(lib.b200gru_gru_cell_fp32_noalloc, c, B, E, H)
        assert rc_a == 0 and rc_b == 0, lib.b200gru_last_error()
        # the fp32 path is checked bit for bit everywhere else in this suite
        assert torch.equal(got, want)
        assert torch.isfinite(got).all()


@pytest.mark.gpu
def test_noalloc_does_not_allocate_device_memory():
    lib = _lib()
    B, E, H = 64, 64, 128
    c = _case(B, E, H, 3)
    _call(lib.b200gru_gru_cell_fp32_noalloc, c, B, E, H)      # set-up call on this device
    out = torch.empty(B, H, device="cuda")
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    rc = lib.b200gru_gru_cell_fp32_noalloc(c["x"].data_ptr(), c["h"].data_ptr(), c["w_ih"].data_ptr(),
                                           c["w_hh"].data_ptr(), c["b_ih"].data_ptr(), c["b_hh"].data_ptr(),
                                           out.data_ptr(), B, E, H, c_void_p(torch.cuda.current_stream().cuda_stream))
    free_during, _ = torch.cuda.mem_get_info()
    torch.cuda.synchronize()
    assert rc == 0
    assert free_during == free_before


def test_noalloc_rejects_oversize_with_a_message():
    """Size check happens before any CUDA call, so this runs without a GPU."""
    lib = _lib()
    B, E, H = 4096, 64, 1024                    # 2*4096*3*1024*4 B = 96 MiB > 4 MiB
    dummy = c_void_p(16)                        # never dereferenced: refused first
    rc = lib.b200gru_gru_cell_fp32_noalloc(dummy, dummy, dummy, dummy, dummy, dummy, dummy, B, E, H, None)
    assert rc < 0
    msg = lib.b200gru_last_error().decode()
    assert "scratch" in msg and "4194304" in msg
