"""GPU parity of the on-device transport-block assembly + TB CRC24A (ldpc_cuda_tb_assemble_device,
ldpc_cuda_dematch_decode_tb_batch) against what pusch_decoder_impl::decode does after its codeblock loop
(pusch_decoder_impl.cpp:156-222): concatenation of the K - crc - fillers leading bits of every codeblock message, CRC24A
over the result, tb_crc_ok = all codeblock flags set and remainder zero (one codeblock: its own flag)."""
import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dec():
    d = pkg.LdpcCuda(device=0, max_codeblocks=4096)
    yield d
    d.close()


# (tbs bits, bg, Qm, layers, nof_ch_symbols): several codeblocks with data bits that are / are not whole bytes, one
# codeblock with CRC16 and with CRC24A, the full-slot 152-codeblock block of BASELINE config 3
TB_CASES = [
    (1277992, 1, 8, 4, 42588 * 4),
    (100000, 1, 6, 2, 30000),   # not a 3GPP size: the last codeblock contributes fewer bits (the min() of :198)
    (100008, 1, 6, 2, 30000),
    (25000, 1, 4, 1, 9000),
    (16136, 1, 4, 1, 6000),
    (9992, 2, 2, 1, 12000),
    (3824, 2, 2, 1, 6000),
    (3000, 2, 2, 1, 6000),
    (4008, 1, 2, 1, 3000),
    (24, 2, 2, 1, 200),
]


def _expected_tb(msgs_bits, data_bits, total_bits):
    allb = np.concatenate([m[:data_bits] for m in msgs_bits])[:total_bits]
    return np.packbits(allb)


def _make_case(rng, tbs, bg, Qm, layers, nsym, out_base):
    """Random transport block laid out as decoded codeblock messages (data bits | garbage where CRC and fillers are)."""
    descs, crc_bits = pkg.segment(tbs, bg, Qm, layers, nsym)
    C = descs.size
    Z, F = int(descs[0]["lifting_size"]), int(descs[0]["nof_filler_bits"])
    K = O.BG_K[bg] * Z
    D = K - int(crc_bits[0]) - F
    total = tbs + (24 if C > 1 else 0)
    assert C * D >= total
    tb_bits = rng.integers(0, 2, tbs).astype(np.uint8)
    if C > 1:
        crc = O.port().crc(O.CRC24A, np.packbits(tb_bits), tbs)
        allb = np.concatenate([tb_bits, np.array([(crc >> (23 - k)) & 1 for k in range(24)], np.uint8)])
    else:
        allb = tb_bits
    kb = (K + 7) // 8
    msgs = []
    for r in range(C):
        m = rng.integers(0, 2, kb * 8).astype(np.uint8)  # garbage behind the data bits must not leak into the block
        chunk = allb[r * D:(r + 1) * D]
        m[:chunk.size] = chunk
        msgs.append(m)
        descs[r]["out_offset"] = out_base + r * kb
    return descs, int(crc_bits[0]), msgs, np.packbits(allb), D, total, kb


def test_tb_assemble_device_matches_the_reference_procedure(dec):
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(77)
    all_descs, tbs_list, msgs_all, want = [], [], [], []
    out_base = tb_off = first = 0
    for tbs, bg, Qm, layers, nsym in TB_CASES:
        descs, cb_crc, msgs, tb_bytes, D, total, kb = _make_case(rng, tbs, bg, Qm, layers, nsym, out_base)
        t = np.zeros(1, pkg.TB_DESC)[0]
        t["first_codeblock"], t["nof_codeblocks"], t["tbs_bits"] = first, descs.size, tbs
        t["nof_cb_crc_bits"], t["tb_offset"] = cb_crc, tb_off
        tbs_list.append(t)
        all_descs.append(descs)
        msgs_all += msgs
        want.append((tb_off, tb_bytes, descs.size))
        out_base += descs.size * kb
        tb_off += total // 8 + 5  # odd spacing: blocks start at unaligned bytes
        first += descs.size
    descs = np.concatenate(all_descs)
    tbs_arr = np.array(tbs_list, pkg.TB_DESC)
    n = descs.size
    out = np.concatenate([np.packbits(m) for m in msgs_all])
    res = np.zeros(n, pkg.CB_RESULT)
    res["crc_ok"] = 1
    # transport block 1: one codeblock failed; block 2: one codeblock passed in an earlier transmission (skip_decode);
    # block 3: a flipped data bit (every codeblock flag set, TB CRC must catch it); block 4: clean
    f1 = int(tbs_arr[1]["first_codeblock"]) + 3
    res["crc_ok"][f1] = 0
    f2 = int(tbs_arr[2]["first_codeblock"]) + 1
    res["crc_ok"][f2] = 0
    descs["skip_decode"][f2] = 1
    f3 = int(tbs_arr[3]["first_codeblock"])
    out[int(descs[f3]["out_offset"]) + 7] ^= 0x10
    out_d = torch.from_numpy(out).to(dev)
    res_d = torch.from_numpy(res.view(np.int32)).to(dev)
    tb_d = torch.full((tb_off + 8,), 0xa5, dtype=torch.uint8, device=dev)
    tbres_d = torch.zeros(2 * tbs_arr.size, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    for _ in range(2):  # twice: the accumulators are reset per call
        dec.tb_assemble_device(tbs_arr, descs, out_d.data_ptr(), res_d.data_ptr(), tb_d.data_ptr(), tbres_d.data_ptr())
    dec.synchronize()
    tb = tb_d.cpu().numpy()
    tbres = tbres_d.cpu().numpy().view(pkg.TB_RESULT)
    for i, (off, tb_bytes, C) in enumerate(want):
        exp = tb_bytes.copy()
        if i == 3:
            exp[7] ^= 0x10
        got = tb[off:off + exp.size]
        assert (got == exp).all(), (i, np.flatnonzero(got != exp)[:8])
        assert (tb[off + exp.size:off + exp.size + 5] == 0xa5).all(), i  # nothing written behind the block
        rem = O.port().crc(O.CRC24A, exp, exp.size * 8)
        assert int(tbres[i]["crc_remainder"]) == rem, (i, hex(int(tbres[i]["crc_remainder"])), hex(rem))
        nfail = 1 if i == 1 else 0
        assert int(tbres[i]["nof_cb_failed"]) == nfail, i
        ok = nfail == 0 and (C == 1 or rem == 0)
        assert int(tbres[i]["tb_crc_ok"]) == int(ok), i
        if C > 1 and i != 3:
            assert rem == 0
    assert int(tbres[3]["tb_crc_ok"]) == 0 and int(tbres[3]["crc_remainder"]) != 0
    assert int(tbres[2]["tb_crc_ok"]) == 1


def test_tb_assemble_rejects_inconsistent_geometry(dec):
    import torch
    descs, crc_bits = pkg.segment(25000, 1, 4, 1, 9000)
    t = np.zeros(1, pkg.TB_DESC)
    t["nof_codeblocks"], t["tbs_bits"], t["nof_cb_crc_bits"] = descs.size, 25000, 24
    buf = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda:0")
    for field, value in [("tbs_bits", 25001), ("nof_codeblocks", descs.size + 1), ("nof_cb_crc_bits", 16),
                         ("tbs_bits", 8 * 25000)]:
        bad = t.copy()
        bad[field] = value
        with pytest.raises(pkg.LdpcCudaError):
            dec.tb_assemble_device(bad, descs, buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), buf.data_ptr())


@pytest.mark.parametrize("tbs,Qm,layers,nsym", [(100008, 6, 2, 30000), (1277992, 8, 4, 42588 * 4)])
def test_dematch_decode_tb_batch_end_to_end(dec, tbs, Qm, layers, nsym):
    """Two transport blocks in one call through the fused host entry: rate-matched AWGN LLRs in, assembled transport
    blocks + verdicts out. Block 0 is clean; block 1 has one codeblock below its decoding threshold."""
    rng = np.random.default_rng(5)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    descs_l, rms, tb_want = [], [], []
    tbs_arr = np.zeros(2, pkg.TB_DESC)
    first = ro = so = oo = 0
    for b in range(2):
        tb, cws, meta = O.build_transport_block(rng, tbs, 1, Qm, layers, nsym)
        descs, crc_bits = pkg.segment(tbs, 1, Qm, layers, nsym)
        C = descs.size
        assert C == len(cws)
        for r in range(C):
            Z, E, F = int(meta[r, 0]), int(meta[r, 2]), int(meta[r, 3])
            assert int(descs[r]["rm_length"]) == E
            tx = O.port().rate_match(cws[r], E, 0, Qm, 0, F)
            bad = b == 1 and r == C // 2
            rms.append(O.awgn_llrs(rng, tx, -6.0 if bad else 9.0))
            d = descs[r]
            d["early_stop"], d["new_data"], d["in_offset"], d["soft_offset"], d["out_offset"] = 1, 1, ro, so, oo
            d["in_length"] = 66 * Z
            ro += E
            so += 66 * 384
            oo += (22 * Z + 7) // 8
        tbs_arr[b]["first_codeblock"], tbs_arr[b]["nof_codeblocks"], tbs_arr[b]["tbs_bits"] = first, C, tbs
        tbs_arr[b]["nof_cb_crc_bits"], tbs_arr[b]["tb_offset"] = int(crc_bits[0]), b * (tbs // 8 + 3)
        first += C
        descs_l.append(descs)
        tb_want.append(tb)
    descs = np.concatenate(descs_l)
    dec.softbuffer_reserve(so)
    out = np.zeros(oo, np.uint8)
    tb_out = np.zeros(2 * (tbs // 8 + 3), np.uint8)
    _, res, tb_out, tbres = dec.dematch_decode_tb_batch(descs, np.concatenate(rms), out, tbs_arr, tb_out)
    C = int(tbs_arr[0]["nof_codeblocks"])
    assert res["crc_ok"][:C].all()
    assert int(tbres[0]["tb_crc_ok"]) == 1 and int(tbres[0]["crc_remainder"]) == 0 and int(tbres[0]["nof_cb_failed"]) == 0
    assert (tb_out[:tbs // 8] == tb_want[0]).all()
    assert int(tbres[1]["tb_crc_ok"]) == 0 and int(tbres[1]["nof_cb_failed"]) == 1
    # the good codeblocks of the failed block are in place all the same
    D = (tbs + 24) // C
    good = tb_out[tbs // 8 + 3:][:(C // 2) * D // 8]
    assert (good == tb_want[1][:good.size]).all()


def test_dematch_decode_tb_batch_pipelined_in_chunks(dec):
    """16 full-slot transport blocks (2432 codeblocks, 21.8 MB of LLRs) in one call: the host entry splits the call into
    overlapping copy / decode chunks, the assembly runs behind the last of them. One block is sent below threshold."""
    rng = np.random.default_rng(16)
    tbs, Qm, layers, nsym, n_tb = 1277992, 8, 4, 42588 * 4, 16
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    tb, cws, meta = O.build_transport_block(rng, tbs, 1, Qm, layers, nsym)
    descs1, crc_bits = pkg.segment(tbs, 1, Qm, layers, nsym)
    C = descs1.size
    tx = [O.port().rate_match(cws[r], int(meta[r, 2]), 0, Qm, 0, int(meta[r, 3])) for r in range(C)]
    tx_all = np.concatenate(tx)
    E = descs1["rm_length"].astype(np.int64)
    cb_off = np.concatenate([[0], np.cumsum(E)[:-1]])
    kb = 22 * 384 // 8
    descs = np.zeros(n_tb * C, pkg.CB_DESC)
    rms = []
    tbs_arr = np.zeros(n_tb, pkg.TB_DESC)
    stride = tbs // 8 + 3
    for b in range(n_tb):
        s = descs[b * C:(b + 1) * C]
        s[:] = descs1
        s["early_stop"], s["new_data"], s["in_length"] = 1, 1, 66 * 384
        s["in_offset"] = b * tx_all.size + cb_off
        s["soft_offset"] = (b * C + np.arange(C)) * 66 * 384
        s["out_offset"] = (b * C + np.arange(C)) * kb
        rms.append(O.awgn_llrs(rng, tx_all, 3.0 if b == 5 else 9.0))
        tbs_arr[b]["first_codeblock"], tbs_arr[b]["nof_codeblocks"], tbs_arr[b]["tbs_bits"] = b * C, C, tbs
        tbs_arr[b]["nof_cb_crc_bits"], tbs_arr[b]["tb_offset"] = 24, b * stride
    dec.softbuffer_reserve(n_tb * C * 66 * 384)
    out = np.zeros(n_tb * C * kb, np.uint8)
    tb_out = np.zeros(n_tb * stride, np.uint8)
    _, res, tb_out, tbres = dec.dematch_decode_tb_batch(descs, np.concatenate(rms), out, tbs_arr, tb_out)
    for b in range(n_tb):
        if b == 5:
            assert int(tbres[b]["tb_crc_ok"]) == 0 and int(tbres[b]["nof_cb_failed"]) > 0
        else:
            assert int(tbres[b]["tb_crc_ok"]) == 1, b
            assert (tb_out[b * stride:b * stride + tbs // 8] == tb).all(), b
            assert int(tbres[b]["crc_remainder"]) == 0
    # the same call with the codeblock messages left on the device (out = NULL): identical blocks and verdicts
    tb_out2 = np.zeros(n_tb * stride, np.uint8)
    _, res2, tb_out2, tbres2 = dec.dematch_decode_tb_batch(descs, np.concatenate(rms), None, tbs_arr, tb_out2)
    assert (res2 == res).all() and (tbres2 == tbres).all()
    ok = np.flatnonzero(tbres["tb_crc_ok"])
    for b in ok:
        assert (tb_out2[b * stride:(b + 1) * stride] == tb_out[b * stride:(b + 1) * stride]).all(), b
    # ... which a codeblock with skip_decode (its message would have to come from the caller) excludes
    d_skip = descs.copy()
    d_skip["skip_decode"][3] = 1
    with pytest.raises(pkg.LdpcCudaError):
        dec.dematch_decode_tb_batch(d_skip, np.concatenate(rms), None, tbs_arr, tb_out2)
