"""ldpc_cuda_tb_assemble_device_noalloc (the allocation-free variant of the transport-block assembly: scratch in a fixed
file-scope device array) against ldpc_cuda_tb_assemble_device on the same inputs, through the C ABI: assembled bytes,
CRC24A remainders, failed-codeblock counts and verdicts must be identical (integer work: tolerance 0), on two calls in a
row that share the array. Collected last on purpose."""
import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()
pytestmark = pytest.mark.gpu


def _workload(rng, cases):
    """Decoded codeblock messages of random transport blocks + descriptors; a failed codeblock, a skipped one and a flipped
    data bit so that every verdict path is taken."""
    all_descs, tbs_list, msgs = [], [], []
    out_base = tb_off = first = 0
    for tbs, bg, Qm, layers, nsym in cases:
        descs, crc_bits = pkg.segment(tbs, bg, Qm, layers, nsym)
        C, Z, F = descs.size, int(descs[0]["lifting_size"]), int(descs[0]["nof_filler_bits"])
        K = O.BG_K[bg] * Z
        D = K - int(crc_bits[0]) - F
        total = tbs + (24 if C > 1 else 0)
        tb_bits = rng.integers(0, 2, tbs).astype(np.uint8)
        if C > 1:
            crc = O.port().crc(O.CRC24A, np.packbits(tb_bits), tbs)
            tb_bits = np.concatenate([tb_bits, np.array([(crc >> (23 - k)) & 1 for k in range(24)], np.uint8)])
        kb = (K + 7) // 8
        for r in range(C):
            m = rng.integers(0, 2, kb * 8).astype(np.uint8)
            chunk = tb_bits[r * D:(r + 1) * D]
            m[:chunk.size] = chunk
            msgs.append(np.packbits(m))
            descs[r]["out_offset"] = out_base + r * kb
        t = np.zeros(1, pkg.TB_DESC)[0]
        t["first_codeblock"], t["nof_codeblocks"], t["tbs_bits"] = first, C, tbs
        t["nof_cb_crc_bits"], t["tb_offset"] = int(crc_bits[0]), tb_off
        tbs_list.append(t)
        all_descs.append(descs)
        out_base += C * kb
        tb_off += total // 8 + 3
        first += C
    descs = np.concatenate(all_descs)
    tbs_arr = np.array(tbs_list, pkg.TB_DESC)
    out = np.concatenate(msgs)
    res = np.zeros(descs.size, pkg.CB_RESULT)
    res["crc_ok"] = 1
    res["crc_ok"][int(tbs_arr[0]["first_codeblock"]) + int(tbs_arr[0]["nof_codeblocks"]) // 2] = 0
    if tbs_arr.size > 1:
        f = int(tbs_arr[1]["first_codeblock"])
        res["crc_ok"][f] = 0
        descs["skip_decode"][f] = 1
    if tbs_arr.size > 2:
        out[int(descs[int(tbs_arr[2]["first_codeblock"])]["out_offset"]) + 5] ^= 0x04
    return tbs_arr, descs, out, res, tb_off + 8


def test_noalloc_tb_assembly_equals_the_allocating_entry_on_two_calls_in_a_row():
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(2024)
    # call 1 is the larger one, so that call 2 runs on top of its left-over records and remainders
    sets = [_workload(rng, [(1277992, 1, 8, 4, 42588 * 4), (100008, 1, 6, 2, 30000), (25000, 1, 4, 1, 9000),
                            (9992, 2, 2, 1, 12000), (3000, 2, 2, 1, 6000)]),
            _workload(rng, [(319784, 1, 8, 1, 42588), (16136, 1, 4, 1, 6000), (3824, 2, 2, 1, 6000), (24, 2, 2, 1, 200)])]
    dec = pkg.LdpcCuda(device=0, max_codeblocks=256)
    try:
        bufs = []
        for tbs_arr, descs, out, res, tb_bytes in sets:
            assert tbs_arr.size <= pkg.ldpc.NOALLOC_MAX_TBS and descs.size <= pkg.ldpc.NOALLOC_MAX_CBS
            bufs.append(dict(out=torch.from_numpy(out).to(dev), res=torch.from_numpy(res.view(np.int32)).to(dev),
                             tb_new=torch.full((tb_bytes,), 0x5a, dtype=torch.uint8, device=dev),
                             tb_old=torch.full((tb_bytes,), 0x5a, dtype=torch.uint8, device=dev),
                             r_new=torch.full((2 * tbs_arr.size,), -1, dtype=torch.int32, device=dev),
                             r_old=torch.full((2 * tbs_arr.size,), -1, dtype=torch.int32, device=dev)))
        torch.cuda.synchronize()
        # two calls in a row on the handle's stream, nothing in between
        for (tbs_arr, descs, _, _, _), b in zip(sets, bufs):
            dec.tb_assemble_device_noalloc(tbs_arr, descs, b["out"].data_ptr(), b["res"].data_ptr(),
                                           b["tb_new"].data_ptr(), b["r_new"].data_ptr())
        dec.synchronize()
        for (tbs_arr, descs, _, _, _), b in zip(sets, bufs):
            dec.tb_assemble_device(tbs_arr, descs, b["out"].data_ptr(), b["res"].data_ptr(), b["tb_old"].data_ptr(),
                                   b["r_old"].data_ptr())
        dec.synchronize()
        for k, ((tbs_arr, _, _, _, _), b) in enumerate(zip(sets, bufs)):
            new, old = b["tb_new"].cpu().numpy(), b["tb_old"].cpu().numpy()
            assert (new == old).all(), (k, np.flatnonzero(new != old)[:8])
            assert (new != 0x5a).any()
            rn = b["r_new"].cpu().numpy().view(pkg.TB_RESULT)
            ro = b["r_old"].cpu().numpy().view(pkg.TB_RESULT)
            for f in ("crc_remainder", "nof_cb_failed", "tb_crc_ok"):
                assert (rn[f] == ro[f]).all(), (k, f, rn[f], ro[f])
            # the workload exercises the verdict paths: a failed codeblock, a skipped one, a corrupted block, clean ones
            assert int(rn["nof_cb_failed"][0]) == 1 and int(rn["tb_crc_ok"][0]) == 0
            assert int(rn["tb_crc_ok"][1]) == 1
            assert int(rn["tb_crc_ok"][2]) == (0 if int(tbs_arr[2]["nof_codeblocks"]) > 1 else 1)
            assert int(rn["tb_crc_ok"][-1]) == 1
    finally:
        dec.close()


def test_noalloc_tb_assembly_rejects_sizes_beyond_its_fixed_scratch():
    import torch
    dec = pkg.LdpcCuda(device=0, max_codeblocks=64)
    try:
        buf = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda:0")
        p = buf.data_ptr()
        q = (p + 8192, p + 16384, p + 32768)  # results, assembled block, verdicts
        descs, crc_bits = pkg.segment(3000, 2, 2, 1, 6000)
        t = np.zeros(1, pkg.TB_DESC)
        t["nof_codeblocks"], t["tbs_bits"], t["nof_cb_crc_bits"] = 1, 3000, int(crc_bits[0])
        dec.tb_assemble_device_noalloc(t, descs, p, *q)  # within the limits: accepted
        too_many_tbs = np.repeat(t, pkg.ldpc.NOALLOC_MAX_TBS + 1)
        too_many_cbs = np.repeat(descs, pkg.ldpc.NOALLOC_MAX_CBS + 1)
        too_long = t.copy()
        too_long["tbs_bits"] = 8 * 1024 * (pkg.ldpc.NOALLOC_MAX_CHUNKS + 1)
        for tbs_arr, d in ((too_many_tbs, descs), (t, too_many_cbs), (too_long, descs)):
            with pytest.raises(pkg.LdpcCudaError) as e:
                dec.tb_assemble_device_noalloc(tbs_arr, d, p, *q)
            assert e.value.status == -3
        dec.synchronize()
    finally:
        dec.close()
