"""Several devices from one process (ldpc_cuda_pool_*): a call's codeblocks go round-robin over the devices
(i -> device i mod G), transport blocks stay whole and retransmissions go to the device that holds their soft buffers;
nothing is exchanged between devices. Results must equal what ONE device computes for the same call, bit for bit (the
single-device entries are compared with the oracle elsewhere). The interleaved-share path of the host entry (pitched copies
out of / into the caller's buffers, other shares' bytes untouched) is exercised on one device too; the tests that need two
devices skip on a smaller box."""
import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()
pytestmark = pytest.mark.gpu

NDEV = pkg.device_count()


def _decode_only_batch(rng, n, bg=1, Z=384, crc=O.CRC24B, esn0=-1.0):
    N, kb = O.BG_N_SHORT[bg] * Z, (O.BG_K[bg] * Z + 7) // 8
    cws = [O.make_codeword(rng, bg, Z, 0, crc)[1] for _ in range(4)]
    llrs = np.stack([O.awgn_llrs(rng, cws[i % 4], esn0 if i % 5 else esn0 - 4.0) for i in range(n)])
    return pkg.make_descs(n, bg, Z, crc_poly=crc, early_stop=1), llrs, kb


def test_an_interleaved_share_reads_and_writes_only_its_own_codeblocks():
    """What a device of a G-device pool does: every G-th codeblock of the caller's buffers, the rest left alone."""
    rng = np.random.default_rng(5)
    n, G = 96, 3
    descs, llrs, kb = _decode_only_batch(rng, n)
    dec = pkg.LdpcCuda(device=0, max_codeblocks=n)
    try:
        ref_out, ref_res = dec.decode_batch(descs, llrs)
        ref_out = ref_out.reshape(n, kb)
        for g in range(G):
            out = np.full((n, kb), 0xc3, np.uint8)
            share = descs[g::G]
            _, res = dec.decode_batch(share, llrs, out.reshape(-1))
            mine = np.arange(g, n, G)
            others = np.setdiff1d(np.arange(n), mine)
            assert (out[mine] == ref_out[mine]).all(), g
            assert (out[others] == 0xc3).all(), g  # bytes of the other shares neither preserved-in nor written back
            assert (res["iterations"] == ref_res["iterations"][mine]).all() and (res["crc_ok"] == ref_res["crc_ok"][mine]).all()
        assert 0 < ref_res["crc_ok"].mean() < 1  # the batch has passing and failing codeblocks
    finally:
        dec.close()


def _pool_devices():
    return list(range(min(NDEV, 8)))


@pytest.mark.parametrize("ndev", [1, 2, 4, 8])
def test_pool_decode_batch_equals_one_device(ndev):
    if ndev > NDEV:
        pytest.skip("needs %d devices, the box has %d" % (ndev, NDEV))
    rng = np.random.default_rng(11 + ndev)
    n = 300 + ndev  # not a multiple of the device count
    descs, llrs, kb = _decode_only_batch(rng, n)
    one = pkg.LdpcCuda(device=0, max_codeblocks=n)
    pool = pkg.LdpcCudaPool(list(range(ndev)), max_codeblocks_per_device=n)
    try:
        want_out, want_res = one.decode_batch(descs, llrs)
        out = np.zeros(n * kb, np.uint8)
        for _ in range(2):  # twice: the second call reuses every device's staging
            _, res = pool.decode_batch(descs, llrs, out)
        assert (out == want_out).all()
        assert (res["iterations"] == want_res["iterations"]).all() and (res["crc_ok"] == want_res["crc_ok"]).all()
        per_dev = pool.codeblocks_per_device()
        assert int(per_dev.sum()) == 2 * n and int(per_dev.max() - per_dev.min()) <= 2  # i mod G
        assert pool.launch_count >= 2 * ndev
    finally:
        pool.close()
        one.close()


@pytest.mark.parametrize("ndev", [1, 2, 8])
def test_pool_decode_batch_device_equals_one_device(ndev):
    """Inputs resident in every device's memory, kernels only: same bits and results as one device."""
    if ndev > NDEV:
        pytest.skip("needs %d devices, the box has %d" % (ndev, NDEV))
    import torch
    rng = np.random.default_rng(40 + ndev)
    n = 130
    descs, llrs, kb = _decode_only_batch(rng, n)
    one = pkg.LdpcCuda(device=0, max_codeblocks=n)
    pool = pkg.LdpcCudaPool(list(range(ndev)), max_codeblocks_per_device=n)
    try:
        want_out, want_res = one.decode_batch(descs, llrs)
        ins = [torch.from_numpy(llrs).to("cuda:%d" % g) for g in range(ndev)]
        outs = [torch.zeros(n * kb, dtype=torch.uint8, device="cuda:%d" % g) for g in range(ndev)]
        ress = [torch.zeros(n, dtype=torch.int32, device="cuda:%d" % g) for g in range(ndev)]
        for g in range(ndev):
            torch.cuda.synchronize(g)
        pool.decode_batch_device(descs, [t.data_ptr() for t in ins], [t.data_ptr() for t in outs],
                                 [t.data_ptr() for t in ress])
        pool.synchronize()
        want_out = want_out.reshape(n, kb)
        for g in range(ndev):
            mine = np.arange(g, n, ndev)
            got = outs[g].cpu().numpy().reshape(n, kb)
            r = ress[g].cpu().numpy().view(pkg.CB_RESULT)[:mine.size]
            assert (got[mine] == want_out[mine]).all(), g
            assert (r["iterations"] == want_res["iterations"][mine]).all() and (r["crc_ok"] == want_res["crc_ok"][mine]).all()
    finally:
        pool.close()
        one.close()


def _tb_slot(rng, geom, rv, new_data, esn0, cws):
    """One slot of transport blocks of one geometry; block t is transmission rv[t] (new_data[t]) of codewords cws[t] (None:
    draw new ones) at Es/N0 esn0[t]. Returns descriptors, rate-matched noisy LLRs, TB descriptors, codewords, sizes."""
    tbs_bits, Qm, layers, nsym = geom
    n_tb = len(rv)
    d0, crc_bits = pkg.segment(tbs_bits, 1, Qm, layers, nsym)
    C = d0.size
    Z, F = int(d0[0]["lifting_size"]), int(d0[0]["nof_filler_bits"])
    kb, Nsoft = (22 * Z + 7) // 8, 66 * Z
    # real transport blocks (payload + CRC24A, CRC24B per codeblock) so that the assembled block passes its own check
    cws = [c if c is not None else O.build_transport_block(rng, tbs_bits, 1, Qm, layers, nsym)[1] for c in cws]
    descs = np.zeros(n_tb * C, pkg.CB_DESC)
    rm, ro = [], 0
    for t in range(n_tb):
        for r in range(C):
            d = descs[t * C + r]
            for f in d0.dtype.names:
                d[f] = d0[r][f]
            E = int(d0[r]["rm_length"])
            rm.append(O.awgn_llrs(rng, O.port().rate_match(cws[t][r], E, rv[t], Qm, 0, F), esn0[t]))
            d["early_stop"], d["new_data"], d["rv"] = 1, new_data[t], rv[t]
            d["in_offset"], d["out_offset"], d["soft_offset"] = ro, (t * C + r) * kb, (t * C + r) * Nsoft
            ro += E
    tbd = np.zeros(n_tb, pkg.TB_DESC)
    tbd["first_codeblock"], tbd["nof_codeblocks"], tbd["tbs_bits"] = np.arange(n_tb) * C, C, tbs_bits
    tbd["nof_cb_crc_bits"] = int(crc_bits[0])
    stride = (tbs_bits // 8 + 3 + 15) // 16 * 16
    tbd["tb_offset"] = np.arange(n_tb, dtype=np.uint64) * np.uint64(stride)
    return descs, np.concatenate(rm), tbd, cws, C, kb, Nsoft, stride


@pytest.mark.parametrize("ndev", [1, 2, 4])
def test_pool_transport_blocks_with_harq_affinity_equal_one_device(ndev):
    """Slot A: 6 new transport blocks, two of them too noisy to decode. Slot B: the rv2 retransmissions of those two
    (new_data = 0) between four new blocks. The pool must send each retransmission to the device that combined its first
    transmission, and every output (messages, codeblock results, assembled blocks, verdicts) must equal the single-device
    run of the same two calls."""
    if ndev > NDEV:
        pytest.skip("needs %d devices, the box has %d" % (ndev, NDEV))
    rng = np.random.default_rng(3)
    n_tb, geom, retx = 6, (23976, 6, 2, 8000), (1, 4)  # 3 codeblocks of BG1 Z=384 per block
    dA, rmA, tbA, cws, C, kb, Nsoft, stride = _tb_slot(rng, geom, [0] * n_tb, [1] * n_tb,
                                                       [-9.0 if t in retx else 5.0 for t in range(n_tb)], [None] * n_tb)
    dB, rmB, tbB, _, _, _, _, _ = _tb_slot(rng, geom, [2 if t in retx else 0 for t in range(n_tb)],
                                           [0 if t in retx else 1 for t in range(n_tb)],
                                           [6.0 if t in retx else 5.0 for t in range(n_tb)],
                                           [cws[t] if t in retx else None for t in range(n_tb)])
    keysA = np.arange(100, 100 + n_tb, dtype=np.uint64)
    keysB = np.array([100 + t if t in retx else 200 + t for t in range(n_tb)], np.uint64)
    arena = n_tb * C * Nsoft

    def run(engine, is_pool):
        outs = []
        engine.softbuffer_reserve(arena)
        for descs, rm, tbd, keys in ((dA, rmA, tbA, keysA), (dB, rmB, tbB, keysB)):
            out = np.zeros(n_tb * C * kb, np.uint8)
            tb_out = np.zeros(n_tb * stride, np.uint8)
            if is_pool:
                _, res, _, tbres = engine.dematch_decode_tb_batch(descs, rm, out, tbd, keys, tb_out)
            else:
                _, res, _, tbres = engine.dematch_decode_tb_batch(descs, rm, out, tbd, tb_out)
            outs.append((out, res.copy(), tb_out, tbres.copy()))
        return outs

    one = pkg.LdpcCuda(device=0, max_codeblocks=n_tb * C)
    pool = pkg.LdpcCudaPool(list(range(ndev)), max_codeblocks_per_device=n_tb * C)
    try:
        want = run(one, False)
        got = run(pool, True)
        for slot, (w, g) in enumerate(zip(want, got)):
            assert (w[0] == g[0]).all(), slot
            assert (w[1]["crc_ok"] == g[1]["crc_ok"]).all() and (w[1]["iterations"] == g[1]["iterations"]).all(), slot
            assert (w[2] == g[2]).all(), slot
            for f in ("tb_crc_ok", "crc_remainder", "nof_cb_failed"):
                assert (w[3][f] == g[3][f]).all(), (slot, f)
        # the scenario does what it says: blocks 1 and 4 fail alone and pass with the retransmission combined
        assert list(want[0][3]["tb_crc_ok"]) == [0 if t in retx else 1 for t in range(n_tb)]
        assert list(want[1][3]["tb_crc_ok"]) == [1] * n_tb
        # affinity: new blocks in turn, retransmissions at the home of their process
        homes = [pool.device_of(int(k)) for k in keysA]
        assert homes == [t % ndev for t in range(n_tb)]
        for t in retx:  # the combined soft buffer of a retransmitted block lives on its home device only
            for g in range(ndev):
                soft = pool.softbuffer_read(g, t * C * Nsoft, Nsoft)
                assert bool(soft.any()) == (g == homes[t]), (t, g)
        assert int(pool.codeblocks_per_device().sum()) == 2 * n_tb * C
    finally:
        pool.close()
        one.close()


def test_two_handles_on_one_device_and_handles_on_two_devices_in_one_process():
    """cudaFuncSetAttribute is per device and the constant tables are uploaded per device: every handle must work whatever
    was created before it."""
    rng = np.random.default_rng(9)
    descs, llrs, kb = _decode_only_batch(rng, 16)
    a = pkg.LdpcCuda(device=0, max_codeblocks=16)
    b = pkg.LdpcCuda(device=0, max_codeblocks=16)
    try:
        want = a.decode_batch(descs, llrs)
        got = b.decode_batch(descs, llrs)
        assert (want[0] == got[0]).all() and (want[1]["iterations"] == got[1]["iterations"]).all()
        if NDEV >= 2:
            c = pkg.LdpcCuda(device=1, max_codeblocks=16)
            try:
                got = c.decode_batch(descs, llrs)
                assert (want[0] == got[0]).all() and (want[1]["iterations"] == got[1]["iterations"]).all()
                # scalar kernel (its own attribute) on the second device as well
                c.set_algorithm(6, 0.75, pkg.SCALE_GENERIC)
                a.set_algorithm(6, 0.75, pkg.SCALE_GENERIC)
                w2, g2 = a.decode_batch(descs, llrs), c.decode_batch(descs, llrs)
                assert (w2[0] == g2[0]).all()
            finally:
                c.close()
    finally:
        a.close()
        b.close()
