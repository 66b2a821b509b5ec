"""Pins the plain-C restatement (oracle/ldpc_oracle.c) against the reference's own compiled code (oracle/_ref).

Runs only where oracle/_ref/libsrsran_ref.so exists (build container, or a GPU box that received the prebuilt .so);
elsewhere the committed fixtures in tests/golden (test_golden.py) play the same role.
The procedures mirror the reference's own unit tests:
  * log_likelihood_ratio_test.cpp:37-64 (LLR arithmetic KATs), hard_decision_test.cpp:35-51,
  * crc_calculator_test.cpp (lut vs clmul vs bit-serial, byte and non-byte lengths),
  * ldpc_enc_dec_test.cpp:249-320 (decode of noiseless codeblocks at several lengths, zero / almost-zero LLRs),
  * ldpc_rm_test.cpp:116-192 (rate match -> dematch round trip over rv, modulation, LBRM, fillers).
"""
import numpy as np
import pytest

import oracles as O

ref = O.ref()
pytestmark = pytest.mark.skipif(ref is None, reason="oracle/_ref not built (no /root/reference on this machine)")

LLR_VALUES = [-127] + list(range(-120, 121)) + [127]


def test_llr_arithmetic_exhaustive():
    p = O.port()
    for a in LLR_VALUES:
        for b in LLR_VALUES:
            assert p.llr_add(a, b) == ref.llr_add(a, b), (a, b)
            assert p.llr_sub(a, b) == ref.llr_sub(a, b), (a, b)
            assert p.llr_promotion_sum(a, b) == ref.llr_promotion_sum(a, b), (a, b)


def test_llr_kats_from_reference_unit_test():
    # tests/unittests/phy/upper/log_likelihood_ratio_test.cpp:37-64
    p = O.port()
    assert p.llr_add(120, 120) == 120 and p.llr_add(-120, -3) == -120
    assert p.llr_add(127, -5) == 127 and p.llr_add(-127, 100) == -127
    assert p.llr_add(127, -127) == 0
    assert p.llr_promotion_sum(120, 120) == 127 and p.llr_promotion_sum(-100, -30) == -127
    assert p.llr_promotion_sum(100, 20) == 120


def test_quantize_matches():
    rng = np.random.default_rng(1)
    xs = np.concatenate([rng.normal(0, 15, 2000), np.arange(-25, 25, 1 / 12.0)]).astype(np.float32)
    q = O.quantize_llr(xs, 20.0)
    for x, v in zip(xs, q):
        assert ref.quantize(x, 20.0) == v, x


def test_hard_decision():
    rng = np.random.default_rng(2)
    for n in [1, 7, 8, 9, 64, 333, 8448]:
        soft = rng.integers(-127, 128, n).astype(np.int8)
        assert (O.port().hard_decision(soft) == ref.hard_decision(soft)).all()


@pytest.mark.parametrize("poly", [O.CRC16, O.CRC24A, O.CRC24B])
def test_crc(poly):
    rng = np.random.default_rng(3)
    for nbits in [1, 5, 8, 16, 23, 24, 25, 31, 40, 100, 1001, 3840, 8448, 8425]:
        data = rng.integers(0, 256, (nbits + 7) // 8).astype(np.uint8)
        if nbits % 8:
            data[-1] &= (0xFF << (8 - nbits % 8)) & 0xFF
        want = ref.crc(poly, data, nbits, "lut")
        assert ref.crc(poly, data, nbits, "auto") == want
        assert O.port().crc(poly, data, nbits) == want, nbits


def test_scale_tables():
    # generic rounding == what the generic decoder applies (checked end to end below); AVX table = avx2_support.h.
    p = O.port()
    gen = [p.scale(m, 0.8, O.SCALE_GENERIC) for m in range(121)]
    avx = [p.scale(m, 0.8, O.SCALE_AVX) for m in range(121)]
    assert gen[1] == 1 and avx[1] == 0 and gen[5] == 4 and avx[5] == 3 and gen[120] == 96 and avx[120] == 95
    assert sum(g != a for g, a in zip(gen, avx)) == 72


@pytest.mark.parametrize("bg", [1, 2])
def test_graphs_all_lifting_sizes(bg):
    p = O.port()
    for Z in O.LIFTING_SIZES:
        pcm, _ = ref.graph(bg, Z)
        for m in range(pcm.shape[0]):
            for n in range(pcm.shape[1]):
                want = -1 if pcm[m, n] == 0xFFFF else int(pcm[m, n])
                assert p.lib.oracle_graph_shift(bg, Z, m, n) == want


@pytest.mark.parametrize("bg", [1, 2])
def test_encoder_all_lifting_sizes(bg):
    rng = np.random.default_rng(4)
    for Z in O.LIFTING_SIZES:
        fillers = int(rng.integers(0, Z)) if Z > 4 else 0
        msg = O.random_message(rng, bg, Z, fillers)
        assert (O.port().encode(bg, Z, msg) == ref.encode(bg, Z, msg)).all(), Z


def _decode_both(bg, Z, llr, **kw):
    out_p, it_p = O.port().decode(bg, Z, llr, **kw)
    out_r, it_r = ref.decode(bg, Z, llr, kind="generic", **kw)
    assert it_p == it_r, (bg, Z, kw, it_p, it_r)
    assert (out_p == out_r).all(), (bg, Z, kw)
    return out_p, it_p


def test_decoder_keeps_padding_bits_of_last_output_byte():
    # bit_buffer::insert (adt/bit_buffer.h:95-140) leaves the bits beyond bg_K * Z in the last byte as the caller had them.
    rng = np.random.default_rng(6)
    for bg, Z in [(2, 9), (1, 2), (2, 3), (1, 6), (2, 15)]:
        assert (O.BG_K[bg] * Z) % 8 != 0
        msg, cw = O.make_codeword(rng, bg, Z)
        llr = O.awgn_llrs(rng, cw, 0.0)
        for out_init in (0x00, 0xFF, 0x5A):
            out, _ = _decode_both(bg, Z, llr, out_init=out_init)
            pad = 8 - (O.BG_K[bg] * Z) % 8
            assert out[-1] & ((1 << pad) - 1) == out_init & ((1 << pad) - 1)


@pytest.mark.parametrize("bg", [1, 2])
def test_decoder_noiseless_all_lifting_sizes(bg):
    # LDPCDecTest, ldpc_enc_dec_test.cpp:249-279: +-10 LLRs, 1 iteration, several shortened lengths.
    rng = np.random.default_rng(5)
    K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
    for Z in O.LIFTING_SIZES:
        msg, cw = O.make_codeword(rng, bg, Z)
        llr = (10 * (1 - 2 * cw.astype(np.int16))).astype(np.int8)
        for nodes in sorted({K + 2, (K + 2 + Ns) // 2, Ns}):
            out, _ = _decode_both(bg, Z, llr[:nodes * Z], max_iters=1)
            assert (out == O.msg_to_packed(msg)).all()


@pytest.mark.parametrize("bg,Z", [(1, 384), (1, 352), (1, 36), (1, 2), (2, 384), (2, 208), (2, 15), (2, 3)])
@pytest.mark.parametrize("crc", [O.CRC_NONE, O.CRC16, O.CRC24A, O.CRC24B])
def test_decoder_noisy(bg, Z, crc):
    rng = np.random.default_rng(100 * bg + Z + crc)
    K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
    for trial, (esn0, iters, fillers) in enumerate([(-1.0, 6, 0), (-3.0, 10, min(12, Z - 1)), (1.0, 1, 0),
                                                    (-4.5, 6, min(28, 2 * Z)), (-2.0, 3, 0)]):
        if crc != O.CRC_NONE and K * Z - fillers <= O.CRC_BITS[crc]:
            continue
        msg, cw = O.make_codeword(rng, bg, Z, fillers, crc)
        llr = O.awgn_llrs(rng, cw, esn0 if bg == 1 else esn0 - 2.0)
        n = Ns * Z if trial % 2 == 0 else (K + 2 + int(rng.integers(2, Ns - K - 1))) * Z
        _decode_both(bg, Z, llr[:n], fillers=fillers, crc=crc, max_iters=iters)


def test_decoder_trailing_zeros_and_non_multiple_lengths():
    rng = np.random.default_rng(7)
    bg, Z = 1, 96
    msg, cw = O.make_codeword(rng, bg, Z, 0, O.CRC24B)
    llr = O.awgn_llrs(rng, cw, -1.0)
    for cut in [30 * Z, 30 * Z + 5, 41 * Z - 1]:
        x = llr.copy()
        x[cut:] = 0  # trailing zeros are trimmed -> fewer layers (ldpc_decoder_impl.cpp:86-114)
        _decode_both(bg, Z, x, crc=O.CRC24B)
        _decode_both(bg, Z, x, crc=O.CRC_NONE)
    for n in [24 * Z + 1, 40 * Z + 17, 66 * Z - 1]:  # fresh decoder object, so no stale tail
        _decode_both(bg, Z, llr[:n], crc=O.CRC24B)


@pytest.mark.parametrize("bg,Z", [(1, 384), (2, 7), (1, 22)])
def test_decoder_zero_and_almost_zero_llrs(bg, Z):
    K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
    zero = np.zeros(Ns * Z, np.int8)
    # LDPCDecTestZeroLLR :296-305
    out, it = _decode_both(bg, Z, zero)
    assert it == 0 and (np.unpackbits(out)[:K * Z] == 1).all()
    # with a CRC the output is left untouched
    out, it = _decode_both(bg, Z, zero, crc=O.CRC16, out_init=0x5A)
    assert it == 0 and (out == 0x5A).all()
    # LDPCDecTestAlmostZeroLLR :307-320
    x = zero.copy()
    for i in range((K + 2) * Z + 2, Ns * Z, 3):
        x[i] = 1 if i % 2 == 0 else -1
    out, it = _decode_both(bg, Z, x)
    assert (np.unpackbits(out)[:K * Z] == 1).all()


def test_decoder_saturated_and_infinite_inputs():
    rng = np.random.default_rng(8)
    bg, Z = 2, 52
    Ns = O.BG_N_SHORT[bg]
    for _ in range(6):
        llr = rng.choice(np.array([-127, -120, -119, -60, -1, 0, 1, 60, 119, 120, 127], np.int8), Ns * Z)
        _decode_both(bg, Z, llr, max_iters=4)
        _decode_both(bg, Z, llr, max_iters=4, crc=O.CRC16)


def test_avx2_scale_mode_agrees_with_reference_avx2():
    if not ref.supports("avx2"):
        pytest.skip("host has no AVX2")
    rng = np.random.default_rng(9)
    agree = total = 0
    for bg, Z in [(1, 384), (2, 160), (1, 56)]:
        for esn0 in [-1.5, -0.5, 1.0]:
            msg, cw = O.make_codeword(rng, bg, Z, 0, O.CRC24A)
            llr = O.awgn_llrs(rng, cw, esn0 - (2.0 if bg == 2 else 0.0))
            out_p, it_p = O.port().decode(bg, Z, llr, crc=O.CRC24A, scale_mode=O.SCALE_AVX)
            out_r, it_r = ref.decode(bg, Z, llr, crc=O.CRC24A, kind="avx2")
            total += 1
            agree += int(it_p == it_r and (out_p == out_r).all())
    # Reported, not gated by the north star; on these inputs the AVX scaling table reproduces avx2 exactly.
    assert agree == total, (agree, total)


RM_CASES = [  # (bg, Z, E, rv, Qm, Nref, fillers) - spread as in ldpc_rm_test.cpp's 25 cases, plus Z=384 shapes
    (1, 16, 208, 0, 1, 0, 0), (1, 16, 300, 0, 2, 0, 12), (1, 16, 2000, 0, 2, 700, 12), (1, 16, 300, 2, 2, 0, 12),
    (1, 16, 1200, 1, 4, 0, 28), (1, 16, 1536, 3, 6, 0, 28), (1, 16, 9240, 3, 8, 700, 12), (2, 12, 600, 0, 4, 0, 28),
    (2, 12, 704, 2, 8, 0, 0), (2, 12, 1200, 1, 6, 430, 12), (1, 384, 8960, 0, 8, 0, 16), (1, 384, 8992, 2, 8, 0, 16),
    (1, 384, 30000, 0, 6, 0, 0), (1, 384, 30000, 3, 2, 16896, 40), (2, 384, 19200, 1, 1, 0, 300),
]


@pytest.mark.parametrize("case", RM_CASES)
def test_rate_match_and_dematch(case):
    bg, Z, E, rv, Qm, Nref, F = case
    rng = np.random.default_rng(hash(case) % (1 << 31))
    msg, cw = O.make_codeword(rng, bg, Z, F)
    rm_p = O.port().rate_match(cw, E, rv, Qm, Nref, F)
    rm_r = ref.rate_match(cw, E, bg, Z, rv, Qm, Nref, F)
    assert (rm_p == rm_r).all()
    llr_in = O.awgn_llrs(rng, rm_p, 2.0)
    N = O.BG_N_SHORT[bg] * Z
    for new_data in (True, False):
        prior = rng.integers(-120, 121, N).astype(np.int8)  # arbitrary prior contents: stale-byte quirks must match
        a, b = prior.copy(), prior.copy()
        O.port().rate_dematch(a, llr_in, new_data, rv, Qm, Nref, F)
        ref.rate_dematch(b, llr_in, new_data, bg, Z, rv, Qm, Nref, F, kind="generic")
        assert (a == b).all(), (case, new_data)
    # ldpc_rm_test.cpp:176-192: dematch(new data, noiseless) then hard-decide == the selected bits
    soft = np.zeros(N, np.int8)
    O.port().rate_dematch(soft, (10 * (1 - 2 * (rm_p & 1).astype(np.int16))).astype(np.int8), True, rv, Qm, Nref, F)
    touched = soft != 0
    assert ((soft[touched] < 0).astype(np.uint8) == (cw[touched] & 1)).all()


def test_harq_chain_rv0231():
    # pusch_decoder_test.cpp:144-175 uses rv sequence {0, 2, 3, 1} into one soft buffer.
    rng = np.random.default_rng(11)
    bg, Z, F, Qm, E = 1, 44, 10, 4, 1400
    msg, cw = O.make_codeword(rng, bg, Z, F, O.CRC24B)
    N = 66 * Z
    a, b = np.zeros(N, np.int8), np.zeros(N, np.int8)
    for i, rv in enumerate([0, 2, 3, 1]):
        rm = O.port().rate_match(cw, E, rv, Qm, 0, F)
        llr = O.awgn_llrs(rng, rm, -2.0)
        O.port().rate_dematch(a, llr, i == 0, rv, Qm, 0, F)
        ref.rate_dematch(b, llr, i == 0, bg, Z, rv, Qm, 0, F)
        assert (a == b).all()
        _decode_both(bg, Z, a, fillers=F, crc=O.CRC24B)


def test_segmenter_geometry():
    p = O.port()
    cases = [(1, 8, 1, 100, 1), (1, 3824, 2, 4000, 1), (1, 3832, 2, 4000, 1), (1, 8424, 4, 6000, 2),
             (1, 8432, 4, 6000, 2), (1, 100000, 6, 30000, 2), (1, 319784, 8, 42588, 1), (1, 400000, 8, 170352, 4),
             (2, 24, 2, 144, 1), (2, 552, 2, 600, 1), (2, 640, 2, 900, 3), (2, 3816, 4, 2400, 2),
             (2, 3840, 4, 2400, 2), (2, 30000, 6, 12000, 4)]
    for bg, tbs, Qm, nsym, nl in cases:
        want = ref.segment_rx(tbs, bg, Qm, nl, nsym)
        got = p.segment_rx(tbs, bg, Qm, nl, nsym)
        assert want is not None and (got == want).all(), (bg, tbs)
    # BASELINE config 3 geometry: beyond the reference's 52-CB cap, port only.
    m = p.segment_rx(1277992, 1, 8, 4, 4 * 42588)
    assert m.shape[0] == 152 and (m[:, 0] == 384).all() and (m[:, 3] == 16).all()
    assert sorted(set(m[:, 2])) == [8960, 8992] and (m[:, 2] == 8960).sum() == 124


# ---- upstream producer: soft demodulator + descrambler ---------------------------------------------
# batch = symbols per call below the width of the reference's AVX2 kernels: only its generic per-symbol code runs
DEMOD_CASES = [(1, False, 0), (1, True, 0), (2, False, 15), (4, False, 7), (6, False, 15), (8, False, 3)]


@pytest.mark.parametrize("mod_order,pi2,batch", DEMOD_CASES)
def test_demodulator_port_equals_reference_generic_path(mod_order, pi2, batch):
    """oracle/demod_oracle.c against demodulation_mapper_impl::demodulate_soft: bit-exact on the generic path; the
    reference's AVX2 kernels (reciprocal-multiply, round-half-even) may differ from its own generic code by one step."""
    ref = O.ref()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(100 + mod_order)
    for noise_var in (0.3, 0.03, 0.003):
        sym, nv = O.random_symbols(rng, mod_order, 60001, noise_var)
        want = ref.demodulate_soft(mod_order, sym, nv, pi2, batch)
        got = O.port().demodulate_soft(mod_order, sym, nv, pi2)
        assert (got == want).all()
        simd = ref.demodulate_soft(mod_order, sym, nv, pi2, 0)
        diff = np.abs(simd.astype(np.int16) - got)
        assert diff.max() <= 1 and (diff != 0).mean() < 1e-4


def test_descrambler_port_equals_reference():
    ref = O.ref()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(9)
    for c_init, n in [(0, 100), (1, 31), (0x7fffffff, 33), ((0x4601 << 15) + 935, 100001), (0x12345678, 8193)]:
        llrs = rng.integers(-127, 128, n).astype(np.int8)
        assert (O.port().descramble(c_init, llrs) == ref.descramble(c_init, llrs)).all()
    # first chips of the Gold sequence for c_init = 0: x2 = 0, so c = x1 advanced by Nc = 1600
    c = O.port().gold_sequence(0, 64)
    x = [1] + [0] * 30
    for n in range(1700):
        x.append(x[n + 3] ^ x[n])
    assert (c == np.array(x[1600:1664], np.uint8)).all()


PLACEHOLDER_CASES = [  # (mod_order, nof_layers, nof_re, placeholder REs)
    (2, 1, 600, [0, 5, 6, 100, 599]), (4, 2, 400, [3, 4, 5, 200, 399]), (6, 1, 1366, [1, 7, 682, 1365]),
    (8, 4, 300, [0, 1, 2, 150, 299]), (8, 1, 3000, [1024, 1025, 2047, 2048]), (1, 1, 500, [0, 2, 4, 250, 498]),
    (6, 3, 911, [0, 455, 910]), (4, 1, 50, [])]


def test_descrambler_with_placeholders_port_equals_reference():
    """oracle_descramble_placeholders against the reference's own pusch_demodulator_impl::descramble
    (pusch_demodulator_impl.cpp:100-152), which the shim calls directly."""
    ref = O.ref()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(10)
    for q, layers, nof_re, res in PLACEHOLDER_CASES:
        rnti, n_id = int(rng.integers(1, 65536)), int(rng.integers(0, 1024))
        llrs = rng.integers(-127, 128, q * layers * nof_re).astype(np.int8)
        want = ref.descramble_placeholders(rnti, n_id, llrs, q, layers, res)
        got = O.port().descramble_placeholders(rnti * 32768 + n_id, llrs, q, layers, res)
        assert (got == want).all(), (q, layers, np.flatnonzero(got != want)[:8])
        if not res:
            assert (got == O.port().descramble(rnti * 32768 + n_id, llrs)).all()
        else:
            assert (got != O.port().descramble(rnti * 32768 + n_id, llrs)).any()
