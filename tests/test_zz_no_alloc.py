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
    try:
        _native.lib()                           # builds the library if the tree is newer
    except (_native.NativeError, RuntimeError, OSError) as e:   # no nvcc and no prebuilt .so here
        pytest.skip(f"native library unavailable: {e}")
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


@pytest.mark.gpu
def test_noalloc_in_place_update():
    """`h_out == h_in` is the way a recurrent caller steps its state: the entry
    point must then update the hidden state in place and still agree with the
    out-of-place result of the allocating twin."""
    lib = _lib()
    B, E, H = 70, 32, 128
    c = _case(B, E, H, 4)
    rc, want = _call(lib.b200gru_gru_cell_fp32, c, B, E, H)
    assert rc == 0
    h = c["h"].clone()
    rc = lib.b200gru_gru_cell_fp32_noalloc(c["x"].data_ptr(), h.data_ptr(), c["w_ih"].data_ptr(),
                                           c["w_hh"].data_ptr(), c["b_ih"].data_ptr(), c["b_hh"].data_ptr(),
                                           h.data_ptr(), B, E, H, c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert rc == 0, lib.b200gru_last_error()
    assert torch.equal(h, want)


@pytest.mark.gpu
def test_noalloc_accepts_rows_that_are_not_16_byte_aligned():
    """A caller that cannot allocate cannot re-align its views either: operands
    offset by one float (and odd inner dimensions) must work, not fault."""
    lib = _lib()
    B, E, H = 19, 5, 32
    g = torch.Generator().manual_seed(5)
    def off(*shape, sc=1.0):                    # a view that starts 4 bytes into its allocation
        n = 1
        for d in shape:
            n *= d
        base = (torch.randn(n + 1, generator=g) * sc).cuda()
        return base[1:].view(*shape)
    c = dict(x=off(B, E), h=off(B, H, sc=0.5), w_ih=off(3 * H, E, sc=0.2), w_hh=off(3 * H, H, sc=0.2),
             b_ih=off(3 * H, sc=0.1), b_hh=off(3 * H, sc=0.1))
    assert c["w_hh"].data_ptr() % 16 != 0
    rc_a, want = _call(lib.b200gru_gru_cell_fp32, c, B, E, H)
    rc_b, got = _call(lib.b200gru_gru_cell_fp32_noalloc, c, B, E, H)
    assert rc_a == 0 and rc_b == 0, lib.b200gru_last_error()
    assert torch.equal(got, want)
    from gru_mpi_cuda_b200.models import oracle
    ref = torch.stack([oracle.gru_cell(c["w_ih"], c["w_hh"], c["b_ih"], c["b_hh"], c["x"][b], c["h"][b])
                       for b in range(B)])
    assert torch.equal(got.cpu(), ref)


@pytest.mark.gpu
def test_noalloc_first_call_is_the_only_one_that_costs_memory():
    """Set-up cost: the first call on a device may make the 4 MiB array (and the
    code around it) resident; it must not cost more than that, and no later
    call may cost anything.  Run in a fresh process so "first" is true."""
    import subprocess, sys, textwrap
    code = textwrap.dedent("""
        import ctypes, sys, torch
        sys.path.insert(0, %r)
        sys.path.insert(0, %r + "/tests")
        from test_zz_no_alloc import _lib, _case
        from ctypes import c_void_p
        lib = _lib()
        B, E, H = 64, 64, 128
        c = _case(B, E, H, 6)
        out = torch.empty(B, H, device="cuda")
        args = (c["x"].data_ptr(), c["h"].data_ptr(), c["w_ih"].data_ptr(), c["w_hh"].data_ptr(),
                c["b_ih"].data_ptr(), c["b_hh"].data_ptr(), out.data_ptr(), B, E, H,
                c_void_p(torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        f0, _ = torch.cuda.mem_get_info()
        assert lib.b200gru_gru_cell_fp32_noalloc(*args) == 0
        torch.cuda.synchronize()
        f1, _ = torch.cuda.mem_get_info()
        assert lib.b200gru_gru_cell_fp32_noalloc(*args) == 0
        f2, _ = torch.cuda.mem_get_info()
        torch.cuda.synchronize()
        print(f0 - f1, f1 - f2)
    """ % ((str(__import__("pathlib").Path(__file__).resolve().parent.parent),) * 2))
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    first, second = map(int, res.stdout.split()[-2:])
    assert second == 0
    # the array is 4 MiB; module images are mapped in 2 MiB granules, so allow a few
    assert 0 <= first <= 32 << 20, first


@pytest.mark.gpu
def test_noalloc_refuses_buffers_of_another_gpu_and_runs_on_device_1():
    """The scratch is per device and the current device's copy is used, so a
    call whose buffers live elsewhere must be refused, and a call made with
    device 1 current must work like one on device 0."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    lib = _lib()
    B, E, H = 40, 32, 64
    c0 = _case(B, E, H, 7)
    rc, want = _call(lib.b200gru_gru_cell_fp32, c0, B, E, H)
    assert rc == 0
    with torch.cuda.device(1):
        out = torch.empty(B, H, device="cuda:1")
        stream = c_void_p(torch.cuda.current_stream().cuda_stream)
        rc = lib.b200gru_gru_cell_fp32_noalloc(c0["x"].data_ptr(), c0["h"].data_ptr(), c0["w_ih"].data_ptr(),
                                               c0["w_hh"].data_ptr(), c0["b_ih"].data_ptr(), c0["b_hh"].data_ptr(),
                                               out.data_ptr(), B, E, H, stream)
        assert rc < 0 and b"device" in lib.b200gru_last_error()
        c1 = {k: v.to("cuda:1") for k, v in c0.items()}
        rc = lib.b200gru_gru_cell_fp32_noalloc(c1["x"].data_ptr(), c1["h"].data_ptr(), c1["w_ih"].data_ptr(),
                                               c1["w_hh"].data_ptr(), c1["b_ih"].data_ptr(), c1["b_hh"].data_ptr(),
                                               out.data_ptr(), B, E, H, stream)
        torch.cuda.synchronize()
        assert rc == 0, lib.b200gru_last_error()
    assert torch.equal(out.cpu(), want.cpu())


def test_noalloc_rejects_oversize_with_a_message():
    """Size check happens before any CUDA call, so this runs without a GPU."""
    lib = _lib()
    B, E, H = 4096, 64, 1024                    # 2*4096*3*1024*4 B = 96 MiB > 4 MiB
    dummy = c_void_p(16)                        # never dereferenced: refused first
    rc = lib.b200gru_gru_cell_fp32_noalloc(dummy, dummy, dummy, dummy, dummy, dummy, dummy, B, E, H, None)
    assert rc < 0
    msg = lib.b200gru_last_error().decode()
    assert "scratch" in msg and "4194304" in msg
