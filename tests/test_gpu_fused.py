"""GPU parity of the batched, fused entry (rate dematch + HARQ combine in the HBM soft-buffer arena -> decode -> CRC early
stop) against the plain-C oracle: what pusch_decoder_impl::decode does per codeblock (pusch_decoder_impl.cpp:156-201),
for a whole batch in one call. Also the pipelined host entry (chunks of copy-in / decode / copy-out that overlap)."""
import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=[0, pkg.FLAG_FORCE_SCALAR], ids=["auto", "scalar"])
def dec(request):
    d = pkg.LdpcCuda(device=0, max_codeblocks=4096, flags=request.param)
    yield d
    d.close()


def _oracle_tx(soft, rm, new_data, bg, Z, rv, Qm, Nref, F, crc, early, max_iters=6):
    """One codeblock of one transmission through the oracle: dematch/combine into soft (in place), then decode."""
    O.port().rate_dematch(soft, rm, new_data, rv, Qm, Nref, F)
    K = O.BG_K[bg]
    if crc and not early:  # decode_cblk without early stop: max iterations, then one CRC check
        bits, _ = O.port().decode(bg, Z, soft, fillers=F, crc=O.CRC_NONE, max_iters=max_iters)
        ok = bool(soft.any()) and O.port().crc(crc, bits, K * Z - F) == 0
        return bits, ok, (max_iters if soft.any() else 0)
    bits, it = O.port().decode(bg, Z, soft, fillers=F, crc=crc, max_iters=max_iters)
    return bits, it > 0, (it if it else (max_iters if soft.any() else 0))


# (bg, Z, Qm, Nref, F, E, crc, esn0)
HARQ_CBS = [
    (1, 384, 8, 0, 16, 8960, O.CRC24B, -2.6),
    (1, 384, 4, 0, 0, 12000, O.CRC24B, -1.5),
    (1, 384, 2, 16896, 40, 9000, O.CRC24A, -3.2),   # limited-buffer rate matching
    (2, 384, 6, 0, 0, 6000, O.CRC24B, -1.0),
    (1, 56, 4, 0, 12, 1400, O.CRC16, 0.0),
    (2, 20, 2, 0, 0, 300, O.CRC16, 3.0),
    (2, 9, 1, 0, 0, 200, O.CRC16, 1.0),
    (1, 208, 6, 0, 28, 4998, O.CRC24B, -4.0),
]


@pytest.mark.parametrize("early", [1, 0])
def test_harq_chain_in_the_arena(dec, early):
    """4 transmissions rv {0,2,3,1} (pusch_decoder_test.cpp:144-175) of a heterogeneous batch; after every transmission the
    soft buffers in HBM, the messages, crc_ok and iterations equal the oracle's. Codeblocks whose CRC has passed are sent
    again with skip_decode (pusch_decoder_impl.cpp:179-184): combined, not decoded, output untouched."""
    rng = np.random.default_rng(31 + early)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    n = len(HARQ_CBS)
    soft_off, out_off, cws, softs = [], [], [], []
    so = oo = 0
    for bg, Z, Qm, Nref, F, E, crc, esn0 in HARQ_CBS:
        soft_off.append(so)
        out_off.append(oo)
        so += (O.BG_N_SHORT[bg] * Z + 255) // 256 * 256
        oo += (O.BG_K[bg] * Z + 7) // 8
        msg, cw = O.make_codeword(rng, bg, Z, F, crc)
        cws.append(cw)
        softs.append(np.zeros(O.BG_N_SHORT[bg] * Z, np.int8))
    dec.softbuffer_reserve(so)
    passed = [False] * n
    out = np.full(oo, 0x5a, np.uint8)
    seen_skip = seen_retx_pass = False
    for t, rv in enumerate([0, 2, 3, 1]):
        descs = np.zeros(n, pkg.CB_DESC)
        rms, ro = [], 0
        want = []
        for i, (bg, Z, Qm, Nref, F, E, crc, esn0) in enumerate(HARQ_CBS):
            tx = O.port().rate_match(cws[i], E, rv, Qm, Nref, F)
            rm = O.awgn_llrs(rng, tx, esn0)
            d = descs[i]
            d["base_graph"], d["lifting_size"], d["nof_filler_bits"], d["crc_poly"] = bg, Z, F, crc
            d["early_stop"], d["rv"], d["mod_order"], d["Nref"], d["new_data"] = early, rv, Qm, Nref, int(t == 0)
            d["rm_length"], d["in_offset"], d["out_offset"], d["soft_offset"] = E, ro, out_off[i], soft_off[i]
            d["in_length"] = O.BG_N_SHORT[bg] * Z
            d["skip_decode"] = int(passed[i])
            rms.append(rm)
            ro += E
            if passed[i]:
                O.port().rate_dematch(softs[i], rm, t == 0, rv, Qm, Nref, F)
                want.append(None)
                seen_skip = True
            else:
                want.append(_oracle_tx(softs[i], rm, t == 0, bg, Z, rv, Qm, Nref, F, crc, early))
        prev_out = out.copy()
        out, res = dec.dematch_decode_batch(descs, np.concatenate(rms), out)
        for i, (bg, Z, Qm, Nref, F, E, crc, esn0) in enumerate(HARQ_CBS):
            N = O.BG_N_SHORT[bg] * Z
            got_soft = dec.softbuffer_read(soft_off[i], N)
            assert (got_soft == softs[i]).all(), (t, i, np.nonzero(got_soft != softs[i])[0][:8])
            kb = (O.BG_K[bg] * Z + 7) // 8
            o = out_off[i]
            if want[i] is None:
                assert (out[o:o + kb] == prev_out[o:o + kb]).all(), (t, i)
                continue
            bits, ok, it = want[i]
            assert int(res["crc_ok"][i]) == int(ok) and int(res["iterations"][i]) == it, (t, i, res[i], ok, it)
            if ok or it:
                nb = O.BG_K[bg] * Z
                assert (np.unpackbits(out[o:o + kb])[:nb] == np.unpackbits(bits)[:nb]).all(), (t, i)
            if ok:
                passed[i] = True
                seen_retx_pass = seen_retx_pass or t > 0
    assert seen_skip and seen_retx_pass and all(passed)


def test_full_slot_transport_block_cfg3(dec):
    """BASELINE config 3: 273 PRB x 13 symbols, 4 layers, 256QAM -> TBS 1 277 992 bits, 152 codeblocks of BG1 Z=384 with
    F=16, E_r = 8960 / 8992, CRC24B early stop. Geometry from ldpc_cuda_segment, one fused call for the TB,
    TB assembled on the host as pusch_decoder_impl.cpp:187-200 does and checked with CRC24A."""
    rng = np.random.default_rng(33)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    tbs, Qm, layers, nsym = 1277992, 8, 4, 42588 * 4
    descs, crc_bits = pkg.segment(tbs, 1, Qm, layers, nsym)
    meta = O.port().segment_rx(tbs, 1, Qm, layers, nsym)
    C = len(descs)
    assert C == 152 and len(meta) == C
    assert set(int(e) for e in descs["rm_length"]) == {8960, 8992}
    assert (descs["lifting_size"] == 384).all() and (descs["nof_filler_bits"] == 16).all() and (crc_bits == 24).all()
    assert (descs["rm_length"] == meta[:, 2]).all()
    # transport block + CRC24A, split into codeblocks with CRC24B and fillers (TS 38.212 5.2.2)
    tb_bits = rng.integers(0, 2, tbs).astype(np.uint8)
    tb_crc = O.port().crc(O.CRC24A, np.packbits(tb_bits), tbs)
    tb_all = np.concatenate([tb_bits, np.array([(tb_crc >> (23 - k)) & 1 for k in range(24)], np.uint8)])
    K, F, Z = 22 * 384, 16, 384
    per_cb = (K - F - 24)
    assert per_cb * C == tb_all.size
    Nsoft = 66 * Z
    dec.softbuffer_reserve(C * Nsoft)
    rms, wants = [], []
    ro = 0
    for r in range(C):
        payload = tb_all[r * per_cb:(r + 1) * per_cb]
        msg = O.attach_crc(np.concatenate([payload, np.zeros(24, np.uint8), np.full(F, 254, np.uint8)]), O.CRC24B, F)
        cw = O.port().encode(1, Z, msg)
        E = int(descs["rm_length"][r])
        tx = O.port().rate_match(cw, E, 0, Qm, 0, F)
        rm = O.awgn_llrs(rng, tx, 9.0 if r % 8 else 6.3)
        d = descs[r]
        d["crc_poly"], d["early_stop"], d["new_data"], d["skip_decode"] = pkg.CRC24B, 1, 1, 0
        d["in_offset"], d["out_offset"], d["soft_offset"] = ro, r * (K // 8), r * Nsoft
        ro += E
        rms.append(rm)
        soft = np.zeros(Nsoft, np.int8)
        wants.append(_oracle_tx(soft, rm, True, 1, Z, 0, Qm, 0, F, O.CRC24B, 1))
    out = np.zeros(C * (K // 8), np.uint8)
    out, res = dec.dematch_decode_batch(descs, np.concatenate(rms), out)
    got_tb = []
    for r, (bits, ok, it) in enumerate(wants):
        assert int(res["crc_ok"][r]) == int(ok) and int(res["iterations"][r]) == it, (r, res[r], ok, it)
        assert (out[r * (K // 8):(r + 1) * (K // 8)] == bits).all(), r
        got_tb.append(np.unpackbits(out[r * (K // 8):(r + 1) * (K // 8)])[:per_cb])
    assert all(w[1] for w in wants), "the test SNR must let every codeblock pass"
    assert len({w[2] for w in wants}) >= 2, "iteration counts should differ between codeblocks"
    got_tb = np.concatenate(got_tb)
    assert (got_tb == tb_all).all()
    assert O.port().crc(O.CRC24A, np.packbits(got_tb), tbs + 24) == 0


def test_host_entry_pipelined_chunks_two_groups(dec):
    """A batch large enough (> 16 MB of LLRs) for the host entry to split it into chunks whose copy-in, kernels and copy-out
    overlap on separate streams, with two (BG, Z) groups: every codeblock must equal the oracle's result of its
    (repeated) input, and output bytes of untouched codeblocks keep the caller's content."""
    rng = np.random.default_rng(34)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    groups = [(1, 384, 24, 42), (2, 384, 10, 60)]  # bg, Z, distinct codewords, repetitions
    llrs, descs, want = [], [], []
    in_off = out_off = 0
    for bg, Z, distinct, reps in groups:
        N, kb = O.BG_N_SHORT[bg] * Z, O.BG_K[bg] * Z // 8
        base = []
        for i in range(distinct):
            msg, cw = O.make_codeword(rng, bg, Z, 0, O.CRC24B)
            x = O.awgn_llrs(rng, cw, (-2.0 if bg == 1 else -5.0) + 0.25 * (i % 8))
            if i == 3:
                x[:] = 0  # all-zero input with a CRC: output untouched (ldpc_decoder_impl.cpp:88-96)
            base.append((x, O.port().decode(bg, Z, x, crc=O.CRC24B, out_init=0x77)))
        for rep in range(reps):
            for i in range(distinct):
                llrs.append(base[i][0])
                want.append(base[i][1])
        n = distinct * reps
        d = pkg.make_descs(n, bg, Z, crc_poly=pkg.CRC24B)
        d["in_offset"] += np.uint64(in_off)
        d["out_offset"] += np.uint64(out_off)
        in_off += n * N
        out_off += n * kb
        descs.append(d)
    descs = np.concatenate(descs)
    assert in_off > (16 << 20)
    out = np.full(out_off, 0x77, np.uint8)
    launches0 = dec.launch_count
    out, res = dec.decode_batch(descs, np.concatenate(llrs), out=out)
    assert dec.launch_count - launches0 > 2, "expected several chunks"
    for i, (bits, it) in enumerate(want):
        o = int(descs["out_offset"][i])
        assert (out[o:o + bits.size] == bits).all(), i
        assert (int(res["iterations"][i]) if res["crc_ok"][i] else 0) == it, i


def test_host_entry_chunks_with_scattered_inputs(dec):
    """The host entry also pipelines when the INPUT layout is not monotone in launch order (new transmissions and
    retransmissions in different regions of the caller's buffer, padding between codeblocks): a chunk then copies the
    contiguous runs of its codeblocks' inputs one by one. Segments of 40 codeblocks placed in a permuted order with gaps;
    results must equal the oracle's for every codeblock."""
    rng = np.random.default_rng(35)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    bg, Z, distinct, reps, seg = 1, 384, 20, 50, 40
    N, kb = O.BG_N_SHORT[bg] * Z, O.BG_K[bg] * Z // 8
    base = []
    for i in range(distinct):
        msg, cw = O.make_codeword(rng, bg, Z, 0, O.CRC24B)
        x = O.awgn_llrs(rng, cw, -2.0 + 0.25 * (i % 8))
        base.append((x, O.port().decode(bg, Z, x, crc=O.CRC24B, out_init=0x33)))
    n = distinct * reps
    stride = N + 64                                  # padding between codeblocks
    nseg = n // seg
    place = rng.permutation(nseg)                    # segment s of the launch order lives at slot place[s]
    seg_bytes = seg * stride + 100000                # and a gap between segments (larger than the run-merging distance)
    buf = np.zeros(nseg * seg_bytes, np.int8)
    descs = pkg.make_descs(n, bg, Z, crc_poly=pkg.CRC24B)
    for i in range(n):
        off = int(place[i // seg]) * seg_bytes + (i % seg) * stride
        descs["in_offset"][i] = off
        buf[off:off + N] = base[i % distinct][0]
    assert n * N > (16 << 20)
    out = np.full(n * kb, 0x33, np.uint8)
    launches0 = dec.launch_count
    out, res = dec.decode_batch(descs, buf, out=out)
    assert dec.launch_count - launches0 > 2, "expected several chunks"
    for i in range(n):
        bits, it = base[i % distinct][1]
        o = int(descs["out_offset"][i])
        assert (out[o:o + bits.size] == bits).all(), i
        assert (int(res["iterations"][i]) if res["crc_ok"][i] else 0) == it, i


@pytest.mark.parametrize("misaligned", [False, True])
def test_fused_prologue_dematch_all_cases_with_stale_priors(dec, misaligned):
    """Every rate-dematcher case of test_gpu_parity.RM_CASES through the FUSED entry: the arena is pre-filled with arbitrary
    prior contents (so the reference's stale-byte / partial-zeroing behaviour is visible), one call carries a new-data and
    a retransmission codeblock of each case plus a skip_decode twin; soft buffers, bits and results must equal the oracle.
    misaligned: soft buffers and rate-matched inputs at odd byte offsets (the kernel's 16-byte / Qm-byte vector accesses
    must fall back to byte accesses)."""
    from test_gpu_parity import RM_CASES
    rng = np.random.default_rng(35)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    cbs = []
    so, oo, ro = (5, 0, 3) if misaligned else (0, 0, 0)
    for case in RM_CASES:
        bg, Z, E, rv, Qm, Nref, F = case
        N, kb = O.BG_N_SHORT[bg] * Z, (O.BG_K[bg] * Z + 7) // 8
        for new_data, skip in ((1, 0), (0, 0), (0, 1)):
            prior = rng.integers(-120, 121, N).astype(np.int8)
            prior[(O.BG_K[bg] - 2) * Z - F:(O.BG_K[bg] - 2) * Z] = 127
            rm = rng.integers(-120, 121, E).astype(np.int8)
            cbs.append(dict(case=case, new_data=new_data, skip=skip, prior=prior, rm=rm, so=so, oo=oo, ro=ro, N=N, kb=kb))
            so += N + 3 if misaligned else (N + 63) // 64 * 64
            oo += kb
            ro += E + (1 if misaligned else 0)
    dec.softbuffer_reserve(so)
    descs = np.zeros(len(cbs), pkg.CB_DESC)
    for i, c in enumerate(cbs):
        bg, Z, E, rv, Qm, Nref, F = c["case"]
        dec.softbuffer_write(c["so"], c["prior"])
        d = descs[i]
        d["base_graph"], d["lifting_size"], d["nof_filler_bits"], d["crc_poly"] = bg, Z, F, O.CRC16
        d["early_stop"], d["rv"], d["mod_order"], d["Nref"], d["new_data"] = 1, rv, Qm, Nref, c["new_data"]
        d["rm_length"], d["in_offset"], d["out_offset"], d["soft_offset"] = E, c["ro"], c["oo"], c["so"]
        d["in_length"], d["skip_decode"] = c["N"], c["skip"]
    rm_all = np.zeros(ro + 16, np.int8)
    for c in cbs:
        rm_all[c["ro"]:c["ro"] + c["rm"].size] = c["rm"]
    out = np.full(oo, 0xa5, np.uint8)
    out, res = dec.dematch_decode_batch(descs, rm_all, out)
    for i, c in enumerate(cbs):
        bg, Z, E, rv, Qm, Nref, F = c["case"]
        want = c["prior"].copy()
        O.port().rate_dematch(want, c["rm"], c["new_data"], rv, Qm, Nref, F)
        got = dec.softbuffer_read(c["so"], c["N"])
        assert (got == want).all(), (c["case"], c["new_data"], c["skip"], np.nonzero(got != want)[0][:10])
        o = c["oo"]
        if c["skip"]:
            assert (out[o:o + c["kb"]] == 0xa5).all(), i
            continue
        bits, it = O.port().decode(bg, Z, want, fillers=F, crc=O.CRC16, out_init=0xa5)
        assert (int(res["iterations"][i]) if res["crc_ok"][i] else 0) == it, (i, c["case"])
        if want.any():
            assert (out[o:o + c["kb"]] == bits).all(), (i, c["case"])


@pytest.mark.parametrize("bg,Z", [(1, 384), (1, 60), (2, 15)])
def test_no_early_stop_half_beside_an_early_finisher(dec, bg, Z):
    """A codeblock decoded without early stop (CRC checked once, after the last iteration; pusch_decoder_impl.cpp:109-118)
    that shares its thread pair with one that stops early: iterations in between run no check at all, and the check after
    the last one must not see what an earlier check left behind. (Found by tests/test_gpu_sweep.py: crc_ok was reported 0
    for correctly decoded bits when the early finisher stopped on an iteration of the same parity as the last.)"""
    rng = np.random.default_rng(77 + Z)
    K, N = O.BG_K[bg], O.BG_N_SHORT[bg] * Z
    for max_iters in (4, 6, 7):
        dec.set_algorithm(max_iters, 0.8, O.SCALE_GENERIC)
        n = 8
        descs = np.zeros(n, pkg.CB_DESC)
        rm_all, want = [], []
        dec.softbuffer_reserve(n * N)
        for i in range(n):
            crc = [O.CRC16, O.CRC24B][i % 2]
            F = (3 * i) % Z
            msg, cw = O.make_codeword(rng, bg, Z, F, crc)
            rm = O.awgn_llrs(rng, O.port().rate_match(cw, N, 0, 2, 0, F), 1.0 + (i // 2))
            d = descs[i]
            d["base_graph"], d["lifting_size"], d["nof_filler_bits"], d["crc_poly"] = bg, Z, F, crc
            d["early_stop"], d["mod_order"], d["new_data"], d["rm_length"] = i % 2, 2, 1, N  # halves of a pair differ
            d["in_offset"], d["out_offset"], d["soft_offset"], d["in_length"] = i * N, i * ((K * Z + 7) // 8), i * N, N
            rm_all.append(rm)
            soft = np.zeros(N, np.int8)
            want.append(_oracle_tx(soft, rm, 1, bg, Z, 0, 2, 0, F, crc, i % 2, max_iters))
        out, res = dec.dematch_decode_batch(descs, np.concatenate(rm_all), np.zeros(n * ((K * Z + 7) // 8), np.uint8))
        kb = (K * Z + 7) // 8
        stopped_early = 0
        for i, (bits, ok, it) in enumerate(want):
            assert bool(res["crc_ok"][i]) == ok and int(res["iterations"][i]) == it, (max_iters, i, res[i], ok, it)
            assert (out[i * kb:(i + 1) * kb] == bits).all()
            stopped_early += ok and it < max_iters - 1
        assert stopped_early >= 2 and all(w[1] for w in want)


@pytest.mark.parametrize("Qm", [2, 4, 6, 8])
def test_wordwise_scatter_first_transmissions(dec, Qm):
    """First transmissions whose geometry is word aligned (E / Qm, the start rank and the filler range multiples of 4) take
    the word-wise scatter of the fused prologue (four symbols transposed in registers, 32-bit shared-memory stores, both
    codeblocks of a pair at once); the same shapes with one parameter off by one symbol / one filler bit take the byte-wise
    scatter. Pairs mix the two. Soft buffers, bits and results equal the oracle's
    (ldpc_rate_dematcher_impl.cpp:125-254, deinterleave :200-210)."""
    rng = np.random.default_rng(60 + Qm)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    cbs = []
    so = oo = ro = 0
    for bg, Z in [(1, 384), (2, 384), (1, 96), (2, 36), (1, 208)]:
        K, N = O.BG_K[bg] * Z, O.BG_N_SHORT[bg] * Z
        for F, extra in [(0, 0), (16, 0), (16, 1), (6, 0), (0, 2), (16, 0)]:
            E = (int(1.2 * K) // (4 * Qm) + extra) * 4 * Qm if extra != 1 else (int(1.2 * K) // (4 * Qm)) * 4 * Qm + Qm
            msg, cw = O.make_codeword(rng, bg, Z, F, O.CRC24B if K > 100 else O.CRC16)
            crc = O.CRC24B if K > 100 else O.CRC16
            rm = O.awgn_llrs(rng, O.port().rate_match(cw, E, 0, Qm, 0, F), 2.0 + 1.5 * Qm / 2 - (2.5 if bg == 2 else 0))
            cbs.append(dict(bg=bg, Z=Z, N=N, F=F, E=E, crc=crc, rm=rm, so=so, oo=oo, ro=ro, kb=(K + 7) // 8))
            so += N
            oo += (K + 7) // 8
            ro += E
    dec.softbuffer_reserve(so)
    dec.softbuffer_write(0, rng.integers(-120, 121, so).astype(np.int8))  # stale contents must not leak into new data
    descs = np.zeros(len(cbs), pkg.CB_DESC)
    for i, c in enumerate(cbs):
        d = descs[i]
        d["base_graph"], d["lifting_size"], d["nof_filler_bits"], d["crc_poly"] = c["bg"], c["Z"], c["F"], c["crc"]
        d["early_stop"], d["rv"], d["mod_order"], d["Nref"], d["new_data"] = 1, 0, Qm, 0, 1
        d["rm_length"], d["in_offset"], d["out_offset"], d["soft_offset"] = c["E"], c["ro"], c["oo"], c["so"]
        d["in_length"] = c["N"]
    out, res = dec.dematch_decode_batch(descs, np.concatenate([c["rm"] for c in cbs]), np.zeros(oo, np.uint8))
    arena = dec.softbuffer_read(0, so)
    passed = 0
    for i, c in enumerate(cbs):
        soft = np.zeros(c["N"], np.int8)
        bits, ok, it = _oracle_tx(soft, c["rm"], True, c["bg"], c["Z"], 0, Qm, 0, c["F"], c["crc"], 1)
        assert (arena[c["so"]:c["so"] + c["N"]] == soft).all(), (i, c["bg"], c["Z"], c["F"], c["E"])
        assert bool(res["crc_ok"][i]) == ok and int(res["iterations"][i]) == it, (i, res[i], ok, it)
        assert (out[c["oo"]:c["oo"] + c["kb"]] == bits).all(), i
        passed += ok
    assert passed > len(cbs) // 2
