from .gemm import gemm_bf16_tn  # noqa: F401
from .gru_cell import gru_cell_fp32  # noqa: F401
