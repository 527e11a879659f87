"""Pins oracle/lsap.c to the third-party solver the reference calls (scipy.optimize.linear_sum_assignment,
ConditionalDETR/models/matcher.py:86): random, integer-tie, constant and rectangular cases, bit-exact."""
import numpy as np
import pytest
from scipy.optimize import linear_sum_assignment as scipy_lsa

from oracle import lsap


def _same(c):
    a, b = lsap.linear_sum_assignment(c)
    a2, b2 = scipy_lsa(c)
    return np.array_equal(a, a2) and np.array_equal(b, b2)


def test_random_and_ties():
    rng = np.random.default_rng(0)
    for trial in range(1500):
        nr, nc = int(rng.integers(1, 48)), int(rng.integers(1, 48))
        kind = trial % 4
        if kind == 0:
            c = rng.standard_normal((nr, nc))
        elif kind == 1:
            c = rng.integers(0, 3, (nr, nc)).astype(float)
        elif kind == 2:
            c = np.zeros((nr, nc))
        else:
            c = rng.standard_normal((nr, nc)).astype(np.float32).astype(float)
            if nc > 2:
                c[:, 1] = c[:, 0]
            if nr > 2:
                c[1, :] = c[0, :]
        assert _same(c), (trial, nr, nc)


@pytest.mark.parametrize("shape", [(128, 300), (300, 128), (128, 0), (128, 1), (128, 128), (128, 129), (128, 256), (6, 3)])
def test_shapes(shape):
    rng = np.random.default_rng(1)
    assert _same(rng.integers(0, 5, shape).astype(float))
    assert _same(rng.standard_normal(shape))


def test_invalid_entries_raise():
    c = np.ones((4, 3))
    c[1, 1] = np.nan
    with pytest.raises(ValueError):
        lsap.linear_sum_assignment(c)
    c[1, 1] = -np.inf
    with pytest.raises(ValueError):
        lsap.linear_sum_assignment(c)
