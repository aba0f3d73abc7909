from .gemm import gemm_bf16_tn  # noqa: F401
