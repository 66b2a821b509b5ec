"""The product computes the circular-buffer start k0 = floor(c_rv * Ncb / N) * Z with an integer division
(srsran_project_23.5_b200/csrc/ldpc_dematch.cuh, make_dematch_geom); the reference computes the quotient in double and
floors it (lib/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher_impl.cpp:101-102). Exhaustive proof that the two agree:
both base graphs, all 51 lifting sizes, every redundancy version, every circular-buffer length 1 .. N."""
import numpy as np

import oracles as O


def test_integer_k0_equals_the_double_computation():
    checked = 0
    for n_short, factors in ((66, (0, 17, 33, 56)), (50, (0, 13, 25, 43))):
        for Z in O.LIFTING_SIZES:
            N = n_short * Z
            ncb = np.arange(1, N + 1, dtype=np.int64)
            for c in factors:
                as_double = np.floor((float(c) * ncb.astype(np.float64)) / float(N)).astype(np.int64)
                assert (as_double == (c * ncb) // N).all(), (n_short, Z, c)
                checked += ncb.size
    assert checked > 2_000_000
