"""Op-level GPU checks against Oracle A ('Kernel' tier)."""
import pytest
import torch

from gru_mpi_cuda_b200.models import oracle
from gru_mpi_cuda_b200.ops import gru_cell_fp32

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,E,H", [(5, 24, 32), (70, 64, 128), (3, 17, 40)])
def test_gru_cell_fp32_bitwise(B, E, H):
    torch.manual_seed(B * 1000 + H)
    w_ih, w_hh = torch.randn(3 * H, E) * 0.2, torch.randn(3 * H, H) * 0.2
    b_ih, b_hh = torch.randn(3 * H) * 0.1, torch.randn(3 * H) * 0.1
    x, h = torch.randn(B, E), torch.randn(B, H) * 0.5
    got = gru_cell_fp32(*[t.cuda() for t in (x, h, w_ih, w_hh, b_ih, b_hh)]).cpu()
    want = torch.stack([oracle.gru_cell(w_ih, w_hh, b_ih, b_hh, x[i], h[i]) for i in range(B)])
    assert torch.equal(got, want)
    cell = torch.nn.GRUCell(E, H)
    cell.load_state_dict({"weight_ih": w_ih, "weight_hh": w_hh, "bias_ih": b_ih, "bias_hh": b_hh})
    assert torch.allclose(got, cell(x, h).detach(), atol=3e-6)
