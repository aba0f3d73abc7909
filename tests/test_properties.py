"""Property tests (hypothesis) for the host-side contracts."""
import math

import torch
from hypothesis import given, settings, strategies as st

from gru_mpi_cuda_b200.models import oracle
from gru_mpi_cuda_b200.parallel import partition


@settings(max_examples=200, deadline=None)
@given(n=st.integers(0, 100000), world=st.integers(1, 64))
def test_partition_covers_a_prefix_without_overlap(n, world):
    slices = [partition(n, world, r) for r in range(world)]
    jpp = n // world
    assert all(s.jpp == jpp and s.count == jpp for s in slices)
    assert [s.start for s in slices] == [jpp * r for r in range(world)]
    assert slices[-1].end == jpp * world <= n          # the N % size tail is dropped (:628)
    kept = [partition(n, world, r, drop_remainder=False) for r in range(world)]
    assert kept[-1].end == n and sum(s.count for s in kept) == n


@settings(max_examples=200, deadline=None)
@given(p=st.lists(st.floats(0.0, 1.0, width=32), min_size=1, max_size=40), r=st.floats(0.0, 1.0, width=32))
def test_random_select_is_the_first_prefix_sum_above_r(p, r):
    probs = torch.tensor(p, dtype=torch.float32)
    got = oracle.random_select(probs, r)
    psum = torch.tensor(0.0)
    want = len(p) - 1
    for i, v in enumerate(probs):
        psum = psum + v                                  # fp32, index order (:326-330)
        if psum.item() > torch.tensor(r, dtype=torch.float32).item():
            want = i
            break
    assert got == want


@settings(max_examples=100, deadline=None)
@given(x=st.lists(st.floats(-20.0, 20.0, width=32), min_size=2, max_size=16))
def test_shared_math_softmax_is_a_distribution_and_close_to_libm(x):
    t = torch.tensor(x, dtype=torch.float32)
    a, b = oracle.softmax(t), oracle.softmax(t, libm=True)
    assert abs(a.sum().item() - 1.0) < 1e-5
    assert torch.allclose(a, b, rtol=5e-7, atol=1e-12)
    assert torch.allclose(a, torch.softmax(t.double(), 0).float(), rtol=2e-6, atol=1e-12)


def test_shared_exp_accuracy():
    """exact_math.h's exp: softmax([x, 0]) gives p0 / p1 = e^x up to two fp32
    division roundings, so the ratio bounds the error of the shared exp."""
    worst = 0.0
    for x in torch.linspace(-10, 10, 401):
        p = oracle.softmax(torch.stack([x, torch.tensor(0.0)])).double()
        ref = math.exp(float(x))
        worst = max(worst, abs(p[0].item() / p[1].item() - ref) / ref)
    assert worst < 4e-7, worst      # ~3 ulp budget: exp (< 1 ulp) + two divisions
