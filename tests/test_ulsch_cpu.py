"""UL-SCH demultiplexing (TS 38.212 6.2.7), CPU side: the plain-C oracle walk (oracle/ulsch_oracle.c) is pinned against the
reference's own ulsch_demultiplex_impl (demultiplex and get_placeholders, compiled from /root/reference) over random
configurations of the message-information space, and the product's host-side closed form (ldpc_cuda_ulsch_sizes /
ldpc_cuda_ulsch_placeholders, no GPU needed) is checked against the oracle. The device kernel uses that same closed form
per resource element and is compared with the oracle in tests/test_gpu_ulsch.py."""
import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()


def _desc(cfg, G_ack, G_c1, G_c2):
    return pkg.ulsch_desc(cfg["mod_order"], cfg["nof_layers"], cfg["nof_prb"], cfg["start_symbol"], cfg["nof_symbols"],
                          cfg["dmrs_type"], cfg["cdm"], cfg["dmrs_mask"], cfg["rvd"], G_ack, G_c1, G_c2)


def _input_bits(cfg):
    """Soft bits of the codeword: elements of every allocated symbol (DM-RS symbols keep the elements left for data)."""
    bpr = cfg["mod_order"] * cfg["nof_layers"]
    n = 0
    for l in range(cfg["start_symbol"], cfg["start_symbol"] + cfg["nof_symbols"]):
        if (cfg["dmrs_mask"] >> l) & 1:
            n += (12 - cfg["cdm"] * (6 if cfg["dmrs_type"] == 1 else 4)) * cfg["nof_prb"]
        else:
            n += 12 * cfg["nof_prb"]
    return n * bpr


def test_oracle_walk_equals_reference_class():
    ref = O.ref()
    if ref is None:
        pytest.skip("oracle/_ref was not built")
    rng = np.random.default_rng(2)
    checked = 0
    for cfg, O_ack, G_ack, O_c1, G_c1, O_c2, G_c2 in O.ulsch_cases(rng, 400):
        llrs = rng.integers(-120, 121, _input_bits(cfg)).astype(np.int8)
        got = O.port().ulsch_demultiplex(cfg, G_ack, G_c1, G_c2, llrs)
        if got is None:
            continue  # the reference asserts on this draw
        want = ref.ulsch_demultiplex(cfg, llrs, got[0].size, G_ack, G_c1, G_c2)
        for a, b in zip(got, want):
            assert (a == b).all(), cfg
        pl = O.port().ulsch_placeholders(cfg, O_ack, G_ack, O_c1, G_c1, O_c2, G_c2)
        assert pl is not None and (pl == ref.ulsch_placeholders(cfg, O_ack, G_ack, O_c1, G_c1, O_c2, G_c2)).all(), cfg
        checked += 1
    assert checked > 150


def test_host_closed_form_equals_oracle():
    rng = np.random.default_rng(3)
    ok = bad = with_pl = 0
    for cfg, O_ack, G_ack, O_c1, G_c1, O_c2, G_c2 in O.ulsch_cases(rng, 1500):
        n_in = _input_bits(cfg)
        llrs = np.zeros(n_in, np.int8)
        want = O.port().ulsch_demultiplex(cfg, G_ack, G_c1, G_c2, llrs)
        d = _desc(cfg, G_ack, G_c1, G_c2)
        sizes = pkg.ulsch_sizes(d)
        if want is None:
            if G_ack or G_c1 or G_c2:
                assert sizes is None, cfg  # inconsistent in the reference -> refused here
            bad += 1
            continue
        assert sizes == (n_in, want[0].size), (cfg, sizes)
        pl = pkg.ulsch_placeholders(d, O_ack, O_c1, O_c2)
        want_pl = O.port().ulsch_placeholders(cfg, O_ack, G_ack, O_c1, G_c1, O_c2, G_c2)
        assert pl is not None and pl.size == want_pl.size and (pl == want_pl).all(), cfg
        with_pl += pl.size > 0
        ok += 1
    assert ok > 600 and bad > 5 and with_pl > 50
