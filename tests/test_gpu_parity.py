"""GPU parity tests proper: the CUDA path, called through the C ABI, against the plain-C oracle on identical inputs.
Bar: hard bits, CRC pass/fail and iteration counts bit-exact (integer path). Marked gpu; run on the B200 box."""
import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()
pytestmark = pytest.mark.gpu

VARIANTS = [0, pkg.FLAG_FORCE_SCALAR]


@pytest.fixture(scope="module", params=VARIANTS, ids=["auto", "scalar"])
def dec(request):
    d = pkg.LdpcCuda(device=0, max_codeblocks=4096, flags=request.param)
    yield d
    d.close()


def check_decode(dec, bg, Z, llr, fillers=0, crc=O.CRC_NONE, max_iters=6, scale_mode=O.SCALE_GENERIC, out_init=0):
    dec.set_algorithm(max_iters, 0.8, scale_mode)
    want_bits, want_it = O.port().decode(bg, Z, llr, fillers=fillers, crc=crc, max_iters=max_iters,
                                         scale_mode=scale_mode, out_init=out_init)
    out = np.full((O.BG_K[bg] * Z + 7) // 8, out_init, np.uint8)
    got_bits, got_it = dec.decode(bg, Z, llr, fillers=fillers, crc=crc, out=out)
    assert got_it == want_it, (bg, Z, crc, max_iters, got_it, want_it)
    assert (got_bits == want_bits).all(), (bg, Z, crc, max_iters)
    return got_bits, got_it


@pytest.mark.parametrize("bg", [1, 2])
def test_decode_all_lifting_sizes(dec, bg):
    """BASELINE config 2: every (BG, Z); noiseless 1-iteration decode (ldpc_enc_dec_test.cpp:249-279) and a noisy
    CRC early-stop decode at a shortened length with filler bits."""
    rng = np.random.default_rng(10 + bg)
    K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
    general0 = dec.general_kernel_launch_count
    for Z in O.LIFTING_SIZES:
        msg, cw = O.make_codeword(rng, bg, Z)
        llr = (10 * (1 - 2 * cw.astype(np.int16))).astype(np.int8)
        for nodes in (K + 2, Ns):
            bits, _ = check_decode(dec, bg, Z, llr[:nodes * Z], max_iters=1)
            assert (bits == O.msg_to_packed(msg)).all()
        F = min(Z - 1, 12) if Z > 3 else 0
        crc = O.CRC16 if K * Z - F <= 40 else O.CRC24B
        msg, cw = O.make_codeword(rng, bg, Z, F, crc)
        noisy = O.awgn_llrs(rng, cw, -1.0 if bg == 1 else -3.5)
        n = (K + 2 + int(rng.integers(2, Ns - K - 1))) * Z
        check_decode(dec, bg, Z, noisy[:n], fillers=F, crc=crc, max_iters=6)
    if not dec.flags & pkg.FLAG_FORCE_SCALAR:
        assert dec.general_kernel_launch_count == general0, "a (BG, Z) fell back to the general decoder kernel"


@pytest.mark.parametrize("bg,Z", [(1, 384), (1, 208), (2, 384), (2, 30)])
@pytest.mark.parametrize("crc", [O.CRC_NONE, O.CRC16, O.CRC24A, O.CRC24B])
def test_decode_noisy_iterations_and_rates(dec, bg, Z, crc):
    rng = np.random.default_rng(1000 * bg + Z + crc)
    K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
    for esn0, iters, fillers, nodes in [(-1.5, 6, 0, Ns), (-2.5, 10, 28, Ns), (1.5, 1, 0, K + 2), (-4.5, 6, 12, Ns),
                                        (0.5, 3, 0, K + 9), (-0.8, 6, 0, Ns)]:
        msg, cw = O.make_codeword(rng, bg, Z, fillers, crc)
        llr = O.awgn_llrs(rng, cw, esn0 - (2.5 if bg == 2 else 0.0))
        check_decode(dec, bg, Z, llr[:nodes * Z], fillers=fillers, crc=crc, max_iters=iters)


def test_decode_zero_almost_zero_and_untouched_output(dec):
    for bg, Z in [(1, 384), (2, 7), (1, 22)]:
        K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
        zero = np.zeros(Ns * Z, np.int8)
        bits, it = check_decode(dec, bg, Z, zero)  # LDPCDecTestZeroLLR
        assert it == 0 and (np.unpackbits(bits)[:K * Z] == 1).all()
        bits, it = check_decode(dec, bg, Z, zero, crc=O.CRC16, out_init=0x5A)  # output left untouched
        assert it == 0 and (bits == 0x5A).all()
        x = zero.copy()  # LDPCDecTestAlmostZeroLLR
        for i in range((K + 2) * Z + 2, Ns * Z, 3):
            x[i] = 1 if i % 2 == 0 else -1
        bits, _ = check_decode(dec, bg, Z, x)
        assert (np.unpackbits(bits)[:K * Z] == 1).all()


def test_decode_keeps_padding_bits_of_last_output_byte(dec):
    rng = np.random.default_rng(6)
    for bg, Z in [(2, 9), (1, 2), (2, 3), (1, 6), (2, 15)]:
        msg, cw = O.make_codeword(rng, bg, Z)
        llr = O.awgn_llrs(rng, cw, 0.0)
        for out_init in (0x00, 0xFF, 0x5A):
            check_decode(dec, bg, Z, llr, out_init=out_init)


def test_decode_trailing_zeros_and_ragged_lengths(dec):
    rng = np.random.default_rng(7)
    for bg, Z in [(1, 96), (2, 72)]:
        K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
        msg, cw = O.make_codeword(rng, bg, Z, 0, O.CRC24B)
        llr = O.awgn_llrs(rng, cw, -1.0 - (2.5 if bg == 2 else 0))
        for cut in [(K + 8) * Z, (K + 8) * Z + 5, (K + 19) * Z - 1]:
            x = llr.copy()
            x[cut:] = 0
            check_decode(dec, bg, Z, x, crc=O.CRC24B)
            check_decode(dec, bg, Z, x)
        for n in [(K + 2) * Z + 1, (K + 18) * Z + 17, Ns * Z - 1]:
            check_decode(dec, bg, Z, llr[:n], crc=O.CRC24B)


def test_decode_saturated_and_infinite_inputs(dec):
    rng = np.random.default_rng(8)
    vals = np.array([-127, -120, -119, -60, -1, 0, 1, 60, 119, 120, 127], np.int8)
    for bg, Z in [(2, 52), (1, 128), (1, 15)]:
        for _ in range(4):
            llr = rng.choice(vals, O.BG_N_SHORT[bg] * Z)
            check_decode(dec, bg, Z, llr, max_iters=4)
            check_decode(dec, bg, Z, llr, max_iters=4, crc=O.CRC16)


def test_decode_avx_and_neon_scaling_modes(dec):
    rng = np.random.default_rng(9)
    for bg, Z in [(1, 384), (2, 160)]:
        msg, cw = O.make_codeword(rng, bg, Z, 0, O.CRC24A)
        llr = O.awgn_llrs(rng, cw, -1.2 - (2.5 if bg == 2 else 0))
        for mode in (O.SCALE_AVX, O.SCALE_NEON, O.SCALE_GENERIC):
            check_decode(dec, bg, Z, llr, crc=O.CRC24A, scale_mode=mode)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)


@pytest.mark.parametrize("sf,mode", [(0.5, O.SCALE_GENERIC), (0.75, O.SCALE_GENERIC), (0.9, O.SCALE_AVX), (0.95, O.SCALE_NEON),
                                     (0.7, O.SCALE_NEON), (0.8, O.SCALE_NEON)])
def test_decode_other_scaling_factors(dec, sf, mode):
    """ldpc_decoder::configuration::algorithm_details::scaling_factor other than the default 0.8, and the NEON variant's
    rounding (ldpc_decoder_neon.cpp scale): the fast kernel takes the scaled magnitude from a 121-entry table
    (TM_SCALE_LUT) instead of its closed forms for 0.8; batches of pairs, every CRC mode, bit-exact with the oracle."""
    rng = np.random.default_rng(int(sf * 100) + mode)
    for bg, Z, esn0 in [(1, 384, -1.3), (2, 384, -4.0), (1, 208, -1.0), (2, 36, -3.0), (1, 7, 0.0)]:
        dec.set_algorithm(6, sf, mode)
        n = 6
        K, N = O.BG_K[bg] * Z, O.BG_N_SHORT[bg] * Z
        llrs = np.zeros((n, N), np.int8)
        descs = pkg.make_descs(n, bg, Z, crc_poly=O.CRC24B if K > 64 else O.CRC16)
        for i in range(n):
            crc = int(descs[i]["crc_poly"])
            msg, cw = O.make_codeword(rng, bg, Z, 0, crc)
            llrs[i] = O.awgn_llrs(rng, cw, esn0 + 0.3 * i)
            descs[i]["early_stop"] = i % 2
        out, res = dec.decode_batch(descs, llrs.reshape(-1))
        kb = (K + 7) // 8
        for i in range(n):
            crc = int(descs[i]["crc_poly"])
            if i % 2:
                bits, it = O.port().decode(bg, Z, llrs[i], crc=crc, max_iters=6, sf=sf, scale_mode=mode)
                ok, iters = it > 0, (it if it else 6)
            else:
                bits, _ = O.port().decode(bg, Z, llrs[i], crc=O.CRC_NONE, max_iters=6, sf=sf, scale_mode=mode)
                ok, iters = O.port().crc(crc, bits, K) == 0, 6
            assert bool(res["crc_ok"][i]) == ok and int(res["iterations"][i]) == iters, (bg, Z, i, res[i], ok, iters)
            assert (out[i * kb:(i + 1) * kb] == bits).all(), (bg, Z, i)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)


def test_invalid_arguments_return_status(dec):
    llr = np.zeros(66 * 384, np.int8)
    with pytest.raises(pkg.LdpcCudaError):
        dec.decode(1, 385, llr)  # not a lifting size
    with pytest.raises(pkg.LdpcCudaError):
        dec.decode(1, 384, llr[:23 * 384])  # shorter than (bg_K + 2) Z, ldpc_decoder_impl.cpp:78-83
    with pytest.raises(pkg.LdpcCudaError):
        dec.set_algorithm(0, 0.8)  # max_iterations > 0
    with pytest.raises(pkg.LdpcCudaError):
        dec.set_algorithm(6, 1.0)  # 0 < sf < 1
    dec.set_algorithm(6, 0.8)


RM_CASES = [  # (bg, Z, E, rv, Qm, Nref, fillers): ldpc_rm_test.cpp's spread + the cfg-3 shapes + LBRM quirk triggers
    (1, 16, 208, 0, 1, 0, 0), (1, 16, 300, 0, 2, 0, 12), (1, 16, 2000, 0, 2, 700, 12), (1, 16, 300, 2, 2, 0, 12),
    (1, 16, 300, 0, 2, 700, 12), (1, 16, 1200, 1, 4, 0, 28), (1, 16, 1536, 3, 6, 0, 28), (1, 16, 9240, 3, 8, 700, 12),
    (2, 12, 600, 0, 4, 0, 28), (2, 12, 704, 2, 8, 0, 0), (2, 12, 1200, 1, 6, 430, 12), (1, 384, 8960, 0, 8, 0, 16),
    (1, 384, 8992, 2, 8, 0, 16), (1, 384, 30000, 0, 6, 0, 0), (1, 384, 30000, 3, 2, 16896, 40),
    (2, 384, 19200, 1, 1, 0, 300), (1, 16, 0, 1, 2, 0, 4), (1, 16, 4, 3, 2, 0, 0), (1, 16, 1056 - 4, 0, 2, 0, 4),
]


@pytest.mark.parametrize("case", RM_CASES)
def test_rate_dematch_matches_oracle(dec, case):
    bg, Z, E, rv, Qm, Nref, F = case
    rng = np.random.default_rng(abs(hash(case)) % (1 << 31))
    N = O.BG_N_SHORT[bg] * Z
    llr_in = rng.integers(-120, 121, E).astype(np.int8)
    for new_data in (True, False):
        prior = rng.integers(-120, 121, N).astype(np.int8)  # arbitrary prior contents: stale-byte quirks must match
        prior[(O.BG_K[bg] - 2) * Z - F:(O.BG_K[bg] - 2) * Z] = 127
        want, got = prior.copy(), prior.copy()
        O.port().rate_dematch(want, llr_in, new_data, rv, Qm, Nref, F)
        dec.rate_dematch(got, llr_in, new_data, bg, Z, rv, Qm, Nref, F)
        assert (want == got).all(), (case, new_data, np.nonzero(want != got)[0][:10])


def test_batch_heterogeneous_codeblocks(dec):
    """Batched entry: codeblocks of different (BG, Z), lengths, fillers, CRCs and early-stop modes in one call,
    including an all-zero and a saturated codeblock; every codeblock must match the oracle individually."""
    rng = np.random.default_rng(21)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    shapes = [(1, 384), (2, 384), (1, 384), (1, 56), (2, 56), (1, 384), (1, 384), (2, 20), (1, 384), (1, 56), (1, 384)]
    llrs, descs, want = [], [], []
    in_off = out_off = 0
    for i, (bg, Z) in enumerate(shapes):
        K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
        crc = [O.CRC24B, O.CRC_NONE, O.CRC24A, O.CRC16][i % 4]
        F = [0, 12, 5, 0][i % 4] if Z > 20 else 0
        early = 0 if i % 5 == 4 else 1
        msg, cw = O.make_codeword(rng, bg, Z, F, crc if crc else O.CRC24B)
        x = O.awgn_llrs(rng, cw, [-1.2, 0.5, -2.5, 3.0][i % 4] - (2.5 if bg == 2 else 0))
        n = Ns * Z if i % 3 else (K + 2 + int(rng.integers(2, Ns - K - 1))) * Z
        x = x[:n].copy()
        if i == 5:
            x[:] = 0
        if i == 8:
            x = rng.choice(np.array([-127, -120, 0, 1, 120, 127], np.int8), n)
        d = pkg.make_descs(1, bg, Z, in_length=n, nof_filler_bits=F, crc_poly=crc, early_stop=early)
        d["in_offset"], d["out_offset"] = in_off, out_off
        kb = (K * Z + 7) // 8
        in_off += n
        out_off += kb
        llrs.append(x)
        descs.append(d)
        if crc and not early:  # decode without CRC for max iterations, then check once (pusch_decoder_impl.cpp:104-117)
            bits, _ = O.port().decode(bg, Z, x, fillers=F, crc=O.CRC_NONE, out_init=0x33)
            ok = O.port().crc(crc, bits, K * Z - F) == 0 and x.any()
            want.append((bits, 6 if x.any() else 0, ok))
        else:
            bits, it = O.port().decode(bg, Z, x, fillers=F, crc=crc, out_init=0x33)
            want.append((bits, it if it else (6 if x.any() else 0), it > 0))
    descs = np.concatenate(descs)
    out = np.full(out_off, 0x33, np.uint8)
    out, res = dec.decode_batch(descs, np.concatenate(llrs), out=out)
    for i, (bits, it, ok) in enumerate(want):
        o = int(descs["out_offset"][i])
        assert (out[o:o + bits.size] == bits).all(), i
        assert int(res["crc_ok"][i]) == int(ok), i
        assert int(res["iterations"][i]) == it, (i, res[i], it)


@pytest.mark.parametrize("bg,Z", [(1, 384), (2, 384), (1, 112), (2, 36)])
def test_batch_every_layer_count(dec, bg, Z):
    """Every number of layers the input length can select (ldpc_decoder_impl.cpp:105-114), in pairs of unequal lengths:
    the kernel's schedule runs rows two at a time where they share no column, so a layer limit that falls inside such a pair
    (and a pair whose halves stop at different rows) must still give the reference's result. The tensor-memory kernel must
    be the one that ran."""
    rng = np.random.default_rng(300 + 10 * bg + Z)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
    nodes = list(range(K + 2, Ns + 1))
    nodes = nodes + [int(v) for v in rng.permutation(nodes)]  # equal-length pairs first, then mixed ones
    msg, cw = O.make_codeword(rng, bg, Z, 0, O.CRC24B)
    full = O.awgn_llrs(rng, cw, -0.5 - (2.5 if bg == 2 else 0))
    kb = (K * Z + 7) // 8
    descs = pkg.make_descs(len(nodes), bg, Z, crc_poly=pkg.CRC24B, in_stride=Ns * Z)
    llrs = np.zeros((len(nodes), Ns * Z), np.int8)
    for i, n in enumerate(nodes):
        descs["in_length"][i] = n * Z
        llrs[i, :n * Z] = full[:n * Z]
        descs["early_stop"][i] = i % 3 != 0  # some halves run all iterations beside an early finisher
    general0 = dec.general_kernel_launch_count
    out, res = dec.decode_batch(descs, llrs.reshape(-1))
    if not dec.flags & pkg.FLAG_FORCE_SCALAR:
        assert dec.general_kernel_launch_count == general0
    for i, n in enumerate(nodes):
        if descs["early_stop"][i]:
            bits, it = O.port().decode(bg, Z, llrs[i, :n * Z], crc=O.CRC24B)
            want_it, ok = (it if it else 6), it > 0
        else:
            bits, _ = O.port().decode(bg, Z, llrs[i, :n * Z], crc=O.CRC_NONE)
            want_it, ok = 6, O.port().crc(O.CRC24B, bits, K * Z) == 0
        assert (out[i * kb:(i + 1) * kb] == bits).all(), (bg, Z, i, n)
        assert int(res["iterations"][i]) == want_it and int(res["crc_ok"][i]) == int(ok), (bg, Z, i, n, res[i])


def test_batch_many_pairs_iteration_counts(dec):
    """200 BG1 Z=384 codeblocks around the waterfall: iteration counts differ inside pairs."""
    rng = np.random.default_rng(22)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    bg, Z, n = 1, 384, 200
    N = 66 * Z
    llrs = np.zeros((n, N), np.int8)
    for i in range(n):
        msg, cw = O.make_codeword(rng, bg, Z, 0, O.CRC24B)
        llrs[i] = O.awgn_llrs(rng, cw, -2.2 + 1.6 * (i % 7) / 6)
    descs = pkg.make_descs(n, bg, Z, crc_poly=pkg.CRC24B)
    out, res = dec.decode_batch(descs, llrs)
    kb = 22 * Z // 8
    its = []
    for i in range(0, n, 3):
        bits, it = O.port().decode(bg, Z, llrs[i], crc=O.CRC24B)
        assert (out[i * kb:(i + 1) * kb] == bits).all(), i
        assert (int(res["iterations"][i]) if res["crc_ok"][i] else 0) == it, i
        its.append(it)
    assert len(set(its)) >= 3
