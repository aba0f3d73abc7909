"""DP partitioner (namegensf.cu:628-630)."""
import pytest

from gru_mpi_cuda_b200.parallel import partition


def test_even_split():
    s = [partition(64, 8, r) for r in range(8)]
    assert all(x.jpp == 8 for x in s)
    assert [x.start for x in s] == list(range(0, 64, 8))
    assert s[-1].end == 64


def test_remainder_is_dropped_like_the_reference():
    s = [partition(10, 4, r) for r in range(4)]
    assert [x.count for x in s] == [2, 2, 2, 2]
    assert s[-1].end == 8                         # names 8, 9 are never generated


def test_remainder_kept_when_asked():
    assert partition(10, 4, 3, drop_remainder=False).end == 10
    assert partition(10, 4, 2, drop_remainder=False).end == 6


def test_fewer_names_than_ranks():
    assert all(partition(3, 8, r).count == 0 for r in range(8))


def test_bad_rank():
    with pytest.raises(ValueError):
        partition(8, 2, 2)
