"""Latency of ONE synchronous ldpc_cuda_dematch_decode_tb_batch call (host buffers, page-locked) as a function of the
number of transport blocks it carries, with and without the codeblock messages travelling back. This is the call the slot
aggregator (host/slot_aggregator.cpp) and pusch_decoder_cuda issue. Run on the GPU box."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import __graft_entry__ as ge  # noqa: E402
import oracles as O  # noqa: E402

pkg = ge.load_package()


def pinned(a):
    t = torch.from_numpy(a).pin_memory()
    return t.numpy(), t


def main():
    rng = np.random.default_rng(3)
    out = []
    for name, tbs, Qm, layers, nsym, esn0 in (("152-codeblock TB R=0.94", 1277992, 8, 4, 42588 * 4, 9.0),
                                               ("12-codeblock TB R=0.5", 100008, 6, 2, 16536 * 2, 3.5)):
        tb, cws, meta = O.build_transport_block(rng, tbs, 1, Qm, layers, nsym)
        C = len(cws)
        tx = np.concatenate([O.port().rate_match(cws[r], int(meta[r, 2]), 0, Qm, 0, int(meta[r, 3])) for r in range(C)])
        llr1 = O.awgn_llrs(rng, tx, esn0)
        d1, crc_bits = pkg.segment(tbs, 1, Qm, layers, nsym)
        kb = (22 * 384 + 7) // 8
        for ntb in (1, 4, 16):
            dec = pkg.LdpcCuda(device=0, max_codeblocks=ntb * C)
            dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
            dec.softbuffer_reserve(ntb * C * 66 * 384)
            descs = np.zeros(ntb * C, pkg.CB_DESC)
            tbd = np.zeros(ntb, pkg.TB_DESC)
            tb_bytes = (tbs + 24 + 7) // 8
            tb_stride = (tb_bytes + 63) // 64 * 64
            for t in range(ntb):
                descs[t * C:(t + 1) * C] = d1
                for r in range(C):
                    d = descs[t * C + r]
                    d["early_stop"], d["new_data"] = 1, 1
                    d["in_offset"] = int(d1[r]["in_offset"]) + t * llr1.size
                    d["soft_offset"] = (t * C + r) * 66 * 384
                    d["out_offset"] = (t * C + r) * kb
                tbd[t]["first_codeblock"], tbd[t]["nof_codeblocks"], tbd[t]["tbs_bits"] = t * C, C, tbs
                tbd[t]["nof_cb_crc_bits"], tbd[t]["tb_offset"] = int(crc_bits[0]), t * tb_stride
            llrs, _k1 = pinned(np.tile(llr1, ntb))
            cb_out, _k2 = pinned(np.zeros(ntb * C * kb, np.uint8))
            tb_out, _k3 = pinned(np.zeros(ntb * tb_stride, np.uint8))
            res, _k4 = pinned(np.zeros(ntb * C, pkg.CB_RESULT).view(np.uint8))
            tbr, _k5 = pinned(np.zeros(ntb, pkg.TB_RESULT).view(np.uint8))
            descs_p, _k6 = pinned(descs.view(np.uint8))
            tbd_p, _k7 = pinned(tbd.view(np.uint8))
            descs_p, tbd_p = descs_p.view(pkg.CB_DESC), tbd_p.view(pkg.TB_DESC)
            res, tbr = res.view(pkg.CB_RESULT), tbr.view(pkg.TB_RESULT)
            for with_out in (True, False):
                ts = []
                for it in range(30):
                    t0 = time.perf_counter()
                    dec.dematch_decode_tb_batch(descs_p, llrs, cb_out if with_out else None, tbd_p, tb_out, res, tbr)
                    ts.append(time.perf_counter() - t0)
                assert all(int(x["tb_crc_ok"]) for x in tbr), name
                assert (tb_out[:tbs // 8] == tb).all()
                ts = np.sort(np.array(ts[5:]))
                out.append({"case": name, "transport_blocks": ntb, "codeblocks": ntb * C, "messages_back": with_out,
                            "median_us": float(np.median(ts) * 1e6), "min_us": float(ts[0] * 1e6),
                            "gbps": ntb * tbs / float(np.median(ts)) / 1e9, "llr_bytes": int(llrs.size)})
                print(out[-1], file=sys.stderr)
            dec.close()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
