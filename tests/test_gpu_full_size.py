"""Parity at BASELINE.json's full size (cfg 4: 65 536 BG1 Z=384 codeblocks, 1.66 GB of LLRs, one launch) through properties
that do not need the oracle on every codeblock:
  * encode -> AWGN -> decode round trip: every codeblock whose CRC24B passes carries exactly the message that was sent, and
    at the bench's SNR all but a handful pass within 6 iterations (the last 4096 codeblocks are sent below threshold so
    that failures and full-length decodes are part of the batch);
  * independence of batch composition: a slice decoded on its own gives the same bytes, verdicts and iteration counts as the
    same codeblocks inside the full batch (no interference between codeblocks at the full grid size, work-stealing order
    does not matter);
  * fixed 6 iterations (the headline configuration, no CRC) equals early stop wherever early stop ran all 6 iterations, and
    a sample of both is bit-exact against the oracle restatement of ldpc_decoder_impl::decode."""
import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()
pytestmark = pytest.mark.gpu

N_CB, BG, Z = 65536, 1, 384
K, N = 22 * Z, 66 * Z


def test_cfg4_full_size_round_trip_and_batch_independence():
    import torch
    import bench
    dev = torch.device("cuda", 0)
    dec = pkg.LdpcCuda(device=0, max_codeblocks=N_CB)
    try:
        rng = np.random.default_rng(1234)
        n_distinct = 256
        msg_d, cw_d = bench.make_codewords_gpu(torch, pkg, dec, dev, rng, n_distinct, BG, Z, 0)
        sent = np.packbits(msg_d.cpu().numpy(), axis=1)                       # [256, K / 8]
        sigma2_good = 10.0 ** (-(-1.0) / 10.0) / 2.0                          # Es/N0 = -1 dB, as bench.py
        sigma2_bad = 10.0 ** (-(-3.5) / 10.0) / 2.0                           # the last 4096: below threshold, all 6 iterations
        g = torch.Generator(device=dev)
        g.manual_seed(4321)
        llr = torch.empty((N_CB, N), dtype=torch.int8, device=dev)
        cwi = cw_d.to(torch.int8)
        for b in range(0, N_CB, 4096):
            sigma2 = sigma2_bad if b + 4096 == N_CB else sigma2_good
            idx = torch.arange(b, b + 4096, device=dev) % n_distinct
            y = (1.0 - 2.0 * cwi[idx].float()) + torch.randn((4096, N), generator=g, device=dev) * (sigma2 ** 0.5)
            x = torch.clamp(2.0 * y / sigma2, -20.0, 20.0) / 20.0 * 120.0
            llr[b:b + 4096] = (torch.sign(x) * torch.floor(torch.abs(x) + 0.5)).to(torch.int8)
        kb = K // 8
        out = torch.zeros((N_CB, kb), dtype=torch.uint8, device=dev)
        res = torch.zeros(N_CB, dtype=torch.int32, device=dev)
        d_es = pkg.make_descs(N_CB, BG, Z, crc_poly=pkg.CRC24B)
        d_es["early_stop"] = 1
        torch.cuda.synchronize()
        dec.decode_batch_device(d_es, llr.data_ptr(), out.data_ptr(), res.data_ptr())
        dec.synchronize()
        out_h = out.cpu().numpy()
        res_h = res.cpu().numpy().view(pkg.CB_RESULT)
        ok = res_h["crc_ok"] == 1
        assert ok[:N_CB - 4096].mean() > 0.999, ok[:N_CB - 4096].mean()
        assert ok[N_CB - 4096:].mean() < 0.5
        assert res_h["iterations"].max() <= 6 and res_h["iterations"][ok].min() >= 1
        want = sent[np.arange(N_CB) % n_distinct]
        assert (out_h[ok] == want[ok]).all(), "a codeblock passed its CRC with a message that was not sent"
        # a slice on its own == the same codeblocks inside the full batch
        for lo, n in ((0, 2048), (40000, 1025), (N_CB - 4096 - 100, 300), (65535, 1)):
            o2 = torch.zeros((n, kb), dtype=torch.uint8, device=dev)
            r2 = torch.zeros(n, dtype=torch.int32, device=dev)
            d2 = pkg.make_descs(n, BG, Z, crc_poly=pkg.CRC24B)
            d2["early_stop"] = 1
            dec.decode_batch_device(d2, llr[lo:lo + n].data_ptr(), o2.data_ptr(), r2.data_ptr())
            dec.synchronize()
            assert (o2.cpu().numpy() == out_h[lo:lo + n]).all(), lo
            assert (r2.cpu().numpy().view(pkg.CB_RESULT) == res_h[lo:lo + n]).all(), lo
        # fixed 6 iterations, no CRC (headline configuration)
        out6 = torch.zeros((N_CB, kb), dtype=torch.uint8, device=dev)
        res6 = torch.zeros(N_CB, dtype=torch.int32, device=dev)
        d6 = pkg.make_descs(N_CB, BG, Z)
        dec.decode_batch_device(d6, llr.data_ptr(), out6.data_ptr(), res6.data_ptr())
        dec.synchronize()
        out6_h = out6.cpu().numpy()
        ran_all = res_h["iterations"] == 6
        assert ran_all.sum() > 2048
        assert (out6_h[ran_all] == out_h[ran_all]).all()     # same 6 iterations, same hard decisions
        # bit-exact against the oracle on a sample spread over the batch (both modes)
        for i in list(rng.integers(0, N_CB, 12)) + [0, N_CB - 1] + list(np.flatnonzero(ran_all)[:4]):
            x = llr[int(i)].cpu().numpy()
            bits, it = O.port().decode(BG, Z, x, crc=O.CRC24B)
            assert (int(res_h["iterations"][i]) if res_h["crc_ok"][i] else 0) == it, i
            if it:
                assert (out_h[i] == bits).all(), i
            bits6, _ = O.port().decode(BG, Z, x, crc=O.CRC_NONE)
            assert (out6_h[i] == bits6).all(), i
    finally:
        dec.close()


def test_cfg3_cfg5_full_slots_harq_properties():
    """BASELINE configs 3 + 5 at full size through the fused entry (bench.py's pusch_slot workload: 16 cells x 152-codeblock
    full-slot transport blocks per slot, soft buffers in the HBM arena, 12.5 % retransmissions): transport blocks sent
    above threshold pass their CRC24A and come back identical to what was sent; those sent below threshold fail with
    every codeblock running all 6 iterations; their rv 2 retransmissions, soft-combined with the stored rv 0 LLRs, pass
    (HARQ gain), also through the host entry."""
    import torch
    import bench
    dev = torch.device("cuda", 0)
    dec = pkg.LdpcCuda(device=0, max_codeblocks=20000)
    try:
        stream = torch.cuda.ExternalStream(dec.stream, device=dev)
        r = bench.pusch_slot_extra(torch, pkg, dec, dev, stream, 1)
    finally:
        dec.close()
    assert r["crc_pass_rate"] == {"new_good_snr": 1.0, "new_bad_snr": 0.0, "retransmission": 1.0}
    assert r["tb_crc_ok_rate"]["new_good_snr"] == 1.0 and r["tb_crc_ok_rate"]["new_bad_snr"] == 0.0
    assert r["tb_crc_ok_rate"]["slot_b_incl_retransmissions"] == 1.0
    assert r["tb_crc_ok_rate"]["assembled_blocks_identical_to_sent"] is True
    assert r["mean_iterations"]["new_bad_snr"] == 6.0 and r["mean_iterations"]["new_good_snr"] < 3.0
    assert r["e2e"].get("tb_crc_ok_rate_slot_b") == 1.0, r["e2e"]
