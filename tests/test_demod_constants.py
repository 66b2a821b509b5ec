"""Exhaustive proof obligations of the demodulator's constant divisions (csrc/ldpc_demod.cuh, demod_div_const): the
two-operation form q = RN(a * hi + RN(a * lo)), hi = RN(1 / d), lo = RN(1 / d - hi), must equal the correctly rounded a / d
of the reference (log_likelihood_ratio.cpp:87-96 `clipped / range_limit`, demodulation_mapper_intervals.h:36-64
`value / interval_width`) for EVERY float a. It is not true for every divisor, so it is checked for the four the kernel
uses, over all 2^23 significands of one binade (the operations are scale invariant for normal numbers; results that
underflow quantise to 0 / interval 0 or -1 either way). No GPU needed: float32 / float64 arithmetic in numpy, where
x * hi + RN(x * lo) is exact in float64 (<= 50 significant bits) and its conversion to float32 is the single rounding of
the fused multiply-add."""
import numpy as np
import pytest


def _widths():
    """Coarse interval widths as build_demod_tables (csrc/ldpc_api.cu) and demod_oracle.c compute them: 2 * (1 / sqrt(n))."""
    return [np.float32(2.0) * (np.float32(1.0) / np.sqrt(np.float32(n))) for n in (42.0, 170.0)]


@pytest.mark.parametrize("divisor", [np.float32(20.0), np.float32(24.0)] + _widths())
def test_two_operation_division_is_correctly_rounded(divisor):
    d = np.float32(divisor)
    hi = np.float32(1.0) / d
    lo = np.float32(np.float64(1.0) / np.float64(d) - np.float64(hi))
    x = (np.arange(1 << 23, 1 << 24, dtype=np.int64).astype(np.float64) * 2.0 ** -23).astype(np.float32)  # [1, 2)
    want = x / d                                                   # IEEE float32 division
    u1 = x * lo                                                    # float32 multiply
    assert want.dtype == np.float32 and u1.dtype == np.float32
    got = (x.astype(np.float64) * np.float64(hi) + u1.astype(np.float64)).astype(np.float32)
    assert (got == want).all(), np.flatnonzero(got != want)[:8]


def test_finest_axis_bit_uses_exactly_doubled_width():
    """The kernel indexes the finest axis bit with the coarse interval (its table line stored twice): valid because the
    reference's fine width is exactly twice the coarse one (4 d vs 2 d, demodulation_mapper_qam64.cpp / _qam256.cpp interval
    tables), so RN(x / (2 w)) = RN(x / w) / 2 and floor(q / 2) = floor(q) >> 1."""
    for n in (42.0, 170.0):
        dd = np.float32(1.0) / np.sqrt(np.float32(n))
        assert np.float32(4.0) * dd == np.float32(2.0) * (np.float32(2.0) * dd)
    q = np.float32(np.random.default_rng(0).normal(0, 40, 1 << 20))
    assert (np.floor(q * np.float32(0.5)).astype(np.int64) == (np.floor(q).astype(np.int64) >> 1)).all()
