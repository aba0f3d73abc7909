"""Stand-alone device ops exposed for tests and profiling."""
from __future__ import annotations

import ctypes

import torch

from .. import _native


def gemm_bf16_tn(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """`D[M,N] = A[M,K] @ B[N,K]^T` on tcgen05 (bf16 in, fp32 out).

    Runs the same TMA -> swizzled smem -> `tcgen05.mma` -> TMEM -> `tcgen05.ld`
    pipeline as the generation kernels (csrc/gru_tc_sm100.cu) with a plain
    store epilogue.  There is no GEMM on a tensor core in the reference; its
    `matvec` is `/root/reference/namegensf.cu:233-242`.
    """
    assert a.is_cuda and b.is_cuda and a.dtype == b.dtype == torch.bfloat16
    a = a.contiguous()
    b = b.contiguous()
    m, k = a.shape
    n, k2 = b.shape
    assert k == k2
    d = torch.empty(m, n, dtype=torch.float32, device=a.device)
    stream = torch.cuda.current_stream(a.device).cuda_stream
    with torch.cuda.device(a.device):
        _native.check(_native.lib().b200gru_gemm_bf16_tn(a.data_ptr(), b.data_ptr(), d.data_ptr(), m, n, k,
                                                        ctypes.c_void_p(stream)), "gemm_bf16_tn")
    return d
