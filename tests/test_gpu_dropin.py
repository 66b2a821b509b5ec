"""Drop-in tests: the reference's UNMODIFIED interfaces and pusch_decoder_impl, compiled from /root/reference into
oracle/_ref/libdropin.so, run on top of the product's "cuda" factories (srsran_project_23.5_b200/host) and are compared
with the same reference code on its own software factories. They follow the reference's own test procedures:
  ldpc_enc_dec_test.cpp:249-320 (decode through ldpc_decoder::decode, zero / almost-zero LLRs),
  ldpc_rm_test.cpp:176-192      (rate_dematch through ldpc_rate_dematcher::rate_dematch),
  pusch_decoder_test.cpp:60-175 (real rx_softbuffer_pool, 4 HARQ transmissions rv {0,2,3,1}, iteration statistics).
Marked gpu; needs oracle/_ref/libdropin.so (built in the build container, shipped to the GPU box)."""
import ctypes
import os

import numpy as np
import pytest

import oracles as O

pytestmark = pytest.mark.gpu

DROPIN = os.path.join(O.ORACLE_DIR, "_ref", "libdropin.so")


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(DROPIN):
        pytest.skip("oracle/_ref/libdropin.so was not built (no reference tree in the build container)")
    return O.RefOracle(DROPIN)


def test_factory_creates_decoder_and_dematcher(lib):
    h = lib.lib.ref_decoder_create(b"cuda")
    assert h, "create_ldpc_decoder_factory_cuda()->create() returned nullptr on a GPU box"
    lib.lib.ref_decoder_destroy(h)
    h = lib.lib.ref_dematcher_create(b"cuda")
    assert h
    lib.lib.ref_dematcher_destroy(h)


@pytest.mark.parametrize("bg,Z", [(1, 384), (1, 52), (2, 384), (2, 9), (1, 2)])
def test_ldpc_decoder_interface(lib, bg, Z):
    """LDPCDecTest: noiseless +-10 LLRs, 1 iteration, several codeblock lengths -> decoded == message; then noisy
    codewords with every CRC through optional<unsigned> decode(...): value / nullopt, iterations and bits identical."""
    rng = np.random.default_rng(100 * bg + Z)
    K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
    msg, cw = O.make_codeword(rng, bg, Z)
    llr = (10 * (1 - 2 * cw.astype(np.int16))).astype(np.int8)
    for nodes in (K + 2, (K + 2 + Ns) // 2, Ns):
        want, wit = lib.decode(bg, Z, llr[:nodes * Z], max_iters=1, kind="generic")
        got, git = lib.decode(bg, Z, llr[:nodes * Z], max_iters=1, kind="cuda")
        assert (got == want).all() and git == wit
        assert (got == O.msg_to_packed(msg)).all()
    for crc in (O.CRC16, O.CRC24A, O.CRC24B):
        if K * Z <= 48:
            continue
        for esn0 in (-0.5, -2.0, -6.0):
            msg, cw = O.make_codeword(rng, bg, Z, 0, crc)
            x = O.awgn_llrs(rng, cw, esn0 - (2.5 if bg == 2 else 0.0))
            want, wit = lib.decode(bg, Z, x, crc=crc, kind="generic", out_init=0x11)
            got, git = lib.decode(bg, Z, x, crc=crc, kind="cuda", out_init=0x11)
            assert git == wit, (bg, Z, crc, esn0, git, wit)
            assert (got == want).all()
    # LDPCDecTestZeroLLR / AlmostZeroLLR
    zero = np.zeros(Ns * Z, np.int8)
    got, git = lib.decode(bg, Z, zero, kind="cuda")
    assert git == 0 and (np.unpackbits(got)[:K * Z] == 1).all()
    x = zero.copy()
    x[(K + 2) * Z + 2::3] = 1
    want, _ = lib.decode(bg, Z, x, kind="generic")
    got, _ = lib.decode(bg, Z, x, kind="cuda")
    assert (got == want).all()


@pytest.mark.parametrize("case", [(1, 16, 300, 0, 2, 0, 12), (1, 16, 2000, 0, 2, 700, 12), (1, 16, 1536, 3, 6, 0, 28),
                                  (2, 12, 704, 2, 8, 0, 0), (1, 384, 8960, 0, 8, 0, 16), (1, 384, 30000, 3, 2, 16896, 40)])
def test_ldpc_rate_dematcher_interface(lib, case):
    bg, Z, E, rv, Qm, Nref, F = case
    rng = np.random.default_rng(sum(case))
    N = O.BG_N_SHORT[bg] * Z
    llr_in = rng.integers(-120, 121, E).astype(np.int8)
    a, b = np.zeros(N, np.int8), np.zeros(N, np.int8)  # pool memory starts at zero (rx_softbuffer_codeblock_pool.h)
    for new_data in (True, False, False):
        lib.rate_dematch(a, llr_in, new_data, bg, Z, rv, Qm, Nref, F, kind="generic")
        lib.rate_dematch(b, llr_in, new_data, bg, Z, rv, Qm, Nref, F, kind="cuda")
        assert (a == b).all(), (case, new_data)


def _pusch(lib, h, slot, tb_bytes, llrs, bg, rv, Qm, nof_layers, iters, early_stop, new_data, rnti=0x4601, harq=3):
    tb = np.zeros(tb_bytes, np.uint8)
    stats = np.zeros(6, np.uint32)
    llrs = np.ascontiguousarray(llrs, np.int8)
    rc = lib.lib.ref_pusch_decode(h, slot, rnti, harq, tb.ctypes.data_as(ctypes.c_void_p), tb_bytes,
                                  llrs.ctypes.data_as(ctypes.c_void_p), llrs.size, bg, rv, Qm, 0, nof_layers, iters,
                                  int(early_stop), int(new_data), 1, stats.ctypes.data_as(ctypes.c_void_p))
    assert rc == 0
    return tb, stats


# Transport block sizes follow TS 38.214 5.1.3.2 (TBS + 24 is a multiple of 8 C): srsvec::copy_offset of the reference
# (lib/srsvec/bit.cpp:333-358) restarts at bit 0 after its byte-aligned fast path, so pusch_decoder_impl loses the last
# (payload % 8) bits of the first codeblock for any other size - 3GPP sizes never trigger it.
@pytest.mark.parametrize("bg,tbs_bytes,Qm,nsym,esn0", [(1, 4001, 4, 12000, -1.0), (2, 300, 2, 2400, -4.5),
                                                      (1, 20004, 6, 42588, 2.0), (2, 1905, 4, 6000, -3.0)])
@pytest.mark.parametrize("early_stop", [True, False])
def test_pusch_decoder_harq_through_reference_pusch_decoder_impl(lib, bg, tbs_bytes, Qm, nsym, esn0, early_stop):
    """pusch_decoder_test.cpp:144-175 with the reference's pusch_decoder_impl on top of (a) its own generic factories and
    (b) the cuda factories: TB bytes, tb_crc_ok and the iteration statistics must agree for every transmission."""
    rng = np.random.default_rng(tbs_bytes + Qm)
    nof_layers = 1
    tb = rng.integers(0, 256, tbs_bytes).astype(np.uint8)
    segs = lib.segment_tx(tb, bg, Qm, nof_layers, nsym)
    meta = lib.segment_rx(tbs_bytes * 8, bg, Qm, nof_layers, nsym)
    assert segs is not None and meta is not None and len(meta) == len(segs)
    h_sw = lib.lib.ref_pusch_create(b"generic", b"generic", 4, 64, 100)
    h_cu = lib.lib.ref_pusch_create(b"cuda", b"cuda", 4, 64, 100)
    assert h_sw and h_cu
    seen_fail = seen_ok = False
    for i, rv in enumerate([0, 2, 3, 1]):
        parts = []
        for c in range(len(segs)):
            Z, N, E, F = int(meta[c, 0]), int(meta[c, 1]), int(meta[c, 2]), int(meta[c, 3])
            K = O.BG_K[bg] * Z
            cw = lib.encode(bg, Z, segs[c, :K])
            parts.append(lib.rate_match(cw, E, bg, Z, rv, Qm, 0, F))
        llrs = O.awgn_llrs(rng, np.concatenate(parts), esn0)
        want_tb, want_st = _pusch(lib, h_sw, 10 + i, tbs_bytes, llrs, bg, rv, Qm, nof_layers, 6, early_stop, i == 0)
        got_tb, got_st = _pusch(lib, h_cu, 10 + i, tbs_bytes, llrs, bg, rv, Qm, nof_layers, 6, early_stop, i == 0)
        assert (want_st == got_st).all(), (i, rv, want_st, got_st)
        if want_st[0]:
            assert (got_tb == want_tb).all() and (got_tb == tb).all()
            seen_ok = True
            break  # the soft buffer is released on success (uplink_processor_impl.cpp:168-170)
        seen_fail = True
    assert seen_ok
    lib.lib.ref_pusch_destroy(h_sw)
    lib.lib.ref_pusch_destroy(h_cu)


@pytest.mark.parametrize("bg,tbs_bytes,Qm,nsym,esn0", [(1, 4001, 4, 12000, -1.0), (2, 300, 2, 2400, -4.5),
                                                      (1, 20004, 6, 42588, 2.0), (2, 1905, 4, 6000, -3.0)])
@pytest.mark.parametrize("early_stop", [True, False])
@pytest.mark.parametrize("pool", [b"", b"cuda-pool"], ids=["ref-pool", "device-pool"])
def test_batched_pusch_decoder_cuda_matches_reference_pusch_decoder(lib, bg, tbs_bytes, Qm, nsym, esn0, early_stop, pool):
    """The product's pusch_decoder (srsran_project_23.5_b200/host/pusch_decoder_cuda.cpp: ONE fused batched call per TB,
    soft bits resident on the device, CRC flags and messages in the rx_softbuffer) against the reference's
    pusch_decoder_impl on its generic factories, over a HARQ sequence rv {0,2,3,1}. Soft buffers come either from the
    reference's own pool (the decoder then keeps the soft bits in a private device arena keyed by the buffer object) or
    from the product's device-resident pool (host/rx_softbuffer_pool_cuda.cpp: one HBM arena, reference FSM)."""
    rng = np.random.default_rng(tbs_bytes + Qm + 1)
    tb = rng.integers(0, 256, tbs_bytes).astype(np.uint8)
    segs = lib.segment_tx(tb, bg, Qm, 1, nsym)
    meta = lib.segment_rx(tbs_bytes * 8, bg, Qm, 1, nsym)
    h_sw = lib.lib.ref_pusch_create(b"generic", b"generic", 4, 64, 100)
    h_cu = lib.lib.ref_pusch_create(b"cuda-batched", pool, 4, 64, 100)
    assert h_sw and h_cu
    seen_ok = False
    for i, rv in enumerate([0, 2, 3, 1]):
        parts = []
        for c in range(len(segs)):
            Z, N, E, F = int(meta[c, 0]), int(meta[c, 1]), int(meta[c, 2]), int(meta[c, 3])
            cw = lib.encode(bg, Z, segs[c, :O.BG_K[bg] * Z])
            parts.append(lib.rate_match(cw, E, bg, Z, rv, Qm, 0, F))
        llrs = O.awgn_llrs(rng, np.concatenate(parts), esn0)
        want_tb, want_st = _pusch(lib, h_sw, 10 + i, tbs_bytes, llrs, bg, rv, Qm, 1, 6, early_stop, i == 0)
        got_tb, got_st = _pusch(lib, h_cu, 10 + i, tbs_bytes, llrs, bg, rv, Qm, 1, 6, early_stop, i == 0)
        assert (want_st == got_st).all(), (i, rv, want_st, got_st)
        if want_st[0]:
            assert (got_tb == want_tb).all() and (got_tb == tb).all()
            seen_ok = True
            break
    assert seen_ok
    lib.lib.ref_pusch_destroy(h_sw)
    lib.lib.ref_pusch_destroy(h_cu)


@pytest.mark.parametrize("pool", [b"", b"cuda-pool"], ids=["shim-softbuffer", "device-pool"])
def test_batched_pusch_decoder_cuda_full_slot_tb_152_codeblocks(lib, pool):
    """BASELINE config 3 through the pusch_decoder interface: TBS 1 277 992 bits = 152 codeblocks, beyond the reference's
    52-codeblock limit (MAX_NOF_SEGMENTS), so the expectation is the transmitted TB itself. First transmission with a
    third of the codeblocks below their decoding threshold (TB fails, the good codeblocks are remembered), rv2
    retransmission: only the failed codeblocks are decoded again, TB CRC24A passes, bytes equal the transmitted TB."""
    rng = np.random.default_rng(152)
    tbs, Qm, layers, nsym = 1277992, 8, 4, 42588 * 4
    tb, cws, meta = O.build_transport_block(rng, tbs, 1, Qm, layers, nsym)
    C = len(cws)
    assert C == 152
    h = lib.lib.ref_pusch_create(b"cuda-batched", pool, 2, 160, 100)
    assert h
    # the reference decoder itself cannot take this transport block
    h_sw = lib.lib.ref_pusch_create(b"generic", b"generic", 2, 52, 100)
    dummy, st6 = np.zeros(tbs // 8, np.uint8), np.zeros(6, np.uint32)
    z = np.zeros(nsym * Qm, np.int8)
    assert lib.lib.ref_pusch_decode(h_sw, 1, 1, 1, dummy.ctypes.data_as(ctypes.c_void_p), tbs // 8,
                                    z.ctypes.data_as(ctypes.c_void_p), z.size, 1, 0, Qm, 0, layers, 6, 1, 1, 1,
                                    st6.ctypes.data_as(ctypes.c_void_p)) == -2
    lib.lib.ref_pusch_destroy(h_sw)
    bad = np.arange(C) % 3 == 0
    nfail = None
    for i, rv in enumerate([0, 2]):
        parts = []
        for r in range(C):
            Z, N, E, F = int(meta[r, 0]), int(meta[r, 1]), int(meta[r, 2]), int(meta[r, 3])
            tx = O.port().rate_match(cws[r], E, rv, Qm, 0, F)
            parts.append(O.awgn_llrs(rng, tx, 4.0 if bad[r] else 9.0))
        llrs = np.concatenate(parts)
        got_tb, st = _pusch(lib, h, 20 + i, tbs // 8, llrs, 1, rv, Qm, layers, 6, True, i == 0)
        assert st[1] == C
        if i == 0:
            assert st[0] == 0 and st[2] == C, st     # TB fails, every codeblock was decoded
            assert st[4] == 6                        # the bad ones ran all iterations
            nfail = int(bad.sum())
        else:
            assert st[0] == 1, st
            assert st[2] == nfail, (st, nfail)      # only the codeblocks that had failed were decoded again
            assert (got_tb == tb).all()
    lib.lib.ref_pusch_destroy(h)


@pytest.mark.parametrize("mod_order,pi2,batch", [(1, False, 0), (1, True, 0), (2, False, 15), (4, False, 7), (6, False, 15),
                                                 (8, False, 3)])
def test_demodulation_mapper_interface(lib, mod_order, pi2, batch):
    """create_demodulation_mapper_cuda() behind the reference's demodulation_mapper interface against
    demodulation_mapper_impl on the same symbols: identical LLRs to the reference's generic per-symbol code, within one
    step of its AVX2 kernels."""
    rng = np.random.default_rng(40 + mod_order)
    sym, nv = O.random_symbols(rng, mod_order, 50003, 0.02)
    got = lib.demodulate_soft(mod_order, sym, nv, pi2, 0xffffffff)
    want = lib.demodulate_soft(mod_order, sym, nv, pi2, batch)
    assert (got == want).all()
    simd = lib.demodulate_soft(mod_order, sym, nv, pi2, 0)
    assert np.abs(simd.astype(np.int16) - got).max() <= 1


def test_descramble_with_placeholders_interface(lib):
    """descramble_cuda(mapper, out, in, pusch_demodulator::configuration) against the reference's own
    pusch_demodulator_impl::descramble (pusch_demodulator_impl.cpp:100-152) on the same configuration: rnti / n_id,
    modulation, layers and a ulsch_placeholder_list."""
    rng = np.random.default_rng(50)
    for q, layers, nof_re, res in [(2, 1, 600, [0, 5, 6, 100, 599]), (4, 2, 400, [3, 4, 5, 200, 399]),
                                   (6, 1, 1366, [1, 7, 682, 1365]), (8, 4, 300, [0, 1, 2, 150, 299]),
                                   (1, 1, 500, [0, 2, 4, 250, 498]), (8, 2, 40000, list(range(3, 40000, 97))), (4, 1, 50, [])]:
        rnti, n_id = int(rng.integers(1, 65536)), int(rng.integers(0, 1024))
        llrs = rng.integers(-127, 128, q * layers * nof_re).astype(np.int8)
        want = lib.descramble_placeholders(rnti, n_id, llrs, q, layers, res)
        got = lib.descramble_placeholders(rnti, n_id, llrs, q, layers, res, cuda=True)
        assert (got == want).all(), (q, layers, np.flatnonzero(got != want)[:8])


def test_ulsch_demultiplex_interface(lib):
    """ulsch_demultiplex (ulsch_demultiplex.h:37-103) created by create_ulsch_demultiplex_cuda(): demultiplex and
    get_placeholders against the reference's own ulsch_demultiplex_impl on the same configuration, as
    ulsch_demultiplex_test.cpp does with its vectors (every soft bit of the four fields, every placeholder index)."""
    rng = np.random.default_rng(31)
    checked = with_pl = 0
    for cfg, O_ack, G_ack, O_c1, G_c1, O_c2, G_c2 in O.ulsch_cases(rng, 160):
        bpr = cfg["mod_order"] * cfg["nof_layers"]
        n_in = 0
        for l in range(cfg["start_symbol"], cfg["start_symbol"] + cfg["nof_symbols"]):
            dm = (cfg["dmrs_mask"] >> l) & 1
            n_in += (12 - dm * cfg["cdm"] * (6 if cfg["dmrs_type"] == 1 else 4)) * cfg["nof_prb"] * bpr
        llrs = rng.integers(-120, 121, n_in).astype(np.int8)
        port = O.port().ulsch_demultiplex(cfg, G_ack, G_c1, G_c2, llrs)
        if port is None:
            continue  # the reference asserts on this draw (and so does the adaptor)
        want = lib.ulsch_demultiplex(cfg, llrs, port[0].size, G_ack, G_c1, G_c2)
        got = lib.ulsch_demultiplex(cfg, llrs, port[0].size, G_ack, G_c1, G_c2, cuda=True)
        for name, a, b in zip(("sch", "ack", "csi1", "csi2"), got, want):
            assert (a == b).all(), (name, cfg)
        wpl = lib.ulsch_placeholders(cfg, O_ack, G_ack, O_c1, G_c1, O_c2, G_c2)
        gpl = lib.ulsch_placeholders(cfg, O_ack, G_ack, O_c1, G_c1, O_c2, G_c2, cuda=True)
        assert gpl.size == wpl.size and (gpl == wpl).all(), cfg
        with_pl += wpl.size > 0
        checked += 1
    assert checked > 60 and with_pl > 5


def _concurrent_llrs(lib, rng, K, reps, nof_tx, bg, tbs_bytes, Qm, nsym, esn0_first, esn0_more):
    """[K][reps][nof_tx][nsym * Qm] LLRs: every (thread, repetition) its own transport block, transmissions rv 0, 2, ..."""
    meta = lib.segment_rx(tbs_bytes * 8, bg, Qm, 1, nsym)
    llrs = np.zeros((K, reps, nof_tx, nsym * Qm), np.int8)
    tbs = np.zeros((K, reps, tbs_bytes), np.uint8)
    for k in range(K):
        for rep in range(reps):
            tb = rng.integers(0, 256, tbs_bytes).astype(np.uint8)
            tbs[k, rep] = tb
            segs = lib.segment_tx(tb, bg, Qm, 1, nsym)
            cws = [lib.encode(bg, int(meta[c, 0]), segs[c, :O.BG_K[bg] * int(meta[c, 0])]) for c in range(len(segs))]
            for t, rv in enumerate([0, 2, 3, 1][:nof_tx]):
                parts = [lib.rate_match(cws[c], int(meta[c, 2]), bg, int(meta[c, 0]), rv, Qm, 0, int(meta[c, 3]))
                         for c in range(len(segs))]
                # every third block starts below its threshold and needs the retransmission
                esn0 = (esn0_first - 3.0 if (k + rep) % 3 == 0 else esn0_first) if t == 0 else esn0_more
                llrs[k, rep, t] = O.awgn_llrs(rng, np.concatenate(parts), esn0)
    return llrs, tbs


@pytest.mark.parametrize("K", [1, 4, 16])
@pytest.mark.parametrize("pool", ["sw", "cuda-pool"], ids=["ref-pool", "device-pool"])
def test_slot_aggregator_concurrent_callers_match_reference(lib, K, pool):
    """K pusch_decoder objects of one factory on K threads (the reference runs one PUSCH per executor thread,
    uplink_processor_impl.cpp:155-171), sharing one soft-buffer pool, each decoding its own transport blocks with a
    retransmission where the first transmission fails: the product's decoders behind the slot aggregator
    (host/slot_aggregator.h - transport blocks that are in decode() at the same time share one launch) against the
    reference's pusch_decoder_impl on K threads. Transport block bytes, tb_crc_ok and the iteration statistics of every
    call are equal; with K > 1 the aggregator really merged calls."""
    rng = np.random.default_rng(40 + K)
    bg, tbs_bytes, Qm, nsym = 1, 4001, 4, 12000
    reps = 6 if K > 1 else 4
    llrs, tbs = _concurrent_llrs(lib, rng, K, reps, 2, bg, tbs_bytes, Qm, nsym, -1.0, 1.0)
    _, want_tb, want_st, _ = lib.pusch_concurrent("generic", "sw", llrs, tbs_bytes, bg, Qm, 1)
    for blocks in (2, 3):
        _, got_tb, got_st, agg = lib.pusch_concurrent("cuda-aggregated", pool, llrs, tbs_bytes, bg, Qm, 1,
                                                      aggregate_blocks=blocks)
        assert (got_st == want_st).all(), np.argwhere(got_st != want_st)[:5]
        ok = want_st[:, :, :, 0].max(axis=2) == 1
        assert ok.all() and (want_st[:, :, 0, 0] == 0).any()      # every block got through, some only when retransmitted
        assert (got_tb == want_tb).all() and (got_tb == tbs).all()
        calls = int((want_st[:, :, :, 1] != 0xffffffff).sum())
        assert agg[1] == calls and agg[0] <= calls
        if K >= 4:
            assert agg[2] >= 2 and agg[0] < calls, agg            # calls were merged into shared launches
