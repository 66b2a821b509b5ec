"""Transport blocks per second behind the reference's own pusch_decoder interface with K concurrent callers (VERDICT r1
item 6): K decoder objects on K threads, one transport block per decode() call, pageable LLR buffers as the reference's
pusch_processor_impl hands them over. Arms: the product's decoder with one launch per call ("cuda-batched"), behind the
slot aggregator with 2 and 3 staging blocks ("cuda-aggregated"), and the reference's pusch_decoder_impl on its avx512 /
avx2 factories on the same K threads. Needs oracle/_ref/libdropin.so. Prints one JSON object; run on the GPU box:
    python tools/pusch_callers_bench.py > gpurun_out/pusch_callers.json"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import oracles as O  # noqa: E402


def main():
    lib = O.RefOracle(os.path.join(O.ORACLE_DIR, "_ref", "libdropin.so"))
    rng = np.random.default_rng(5)
    out = {"what": "transport blocks per second through pusch_decoder::decode, K concurrent callers, new data, CRC early stop",
           "cases": []}
    # (name, bg, tbs_bits, Qm, layers, channel symbols, Es/N0 dB, reps per thread)
    cases = [("full-slot TB, 152 codeblocks (BASELINE config 3: BG1 Z=384, R=0.94, 256QAM x 4 layers)", 1, 1277992, 8, 4, 42588 * 4,
              9.0, 12, "cuda-pool"),
             ("mid-size TB, 12 codeblocks (BG1 Z=384, R=0.5, 64QAM x 2 layers)", 1, 100008, 6, 2, 16536 * 2, 3.5, 40, "sw")]
    for name, bg, tbs, Qm, layers, nsym, esn0, reps, pool in cases:
        tb, cws, meta = O.build_transport_block(rng, tbs, bg, Qm, layers, nsym)
        C = len(cws)
        tx = np.concatenate([O.port().rate_match(cws[r], int(meta[r, 2]), 0, Qm, 0, int(meta[r, 3])) for r in range(C)])
        variants = np.stack([O.awgn_llrs(rng, tx, esn0) for _ in range(4)])
        case = {"name": name, "codeblocks": C, "tbs_bits": tbs, "llrs_per_tb": int(tx.size), "pool": pool, "arms": {}}
        for K in (1, 4, 16):
            llrs = np.ascontiguousarray(variants[(np.arange(K * reps) % 4)].reshape(K, reps, 1, tx.size))
            arms = [("cuda-batched", 3), ("cuda-aggregated", 2), ("cuda-aggregated", 3)]
            if C <= 52:
                arms += [("avx512", 0), ("avx2", 0)]
            for arm, blocks in arms:
                if arm.startswith("avx") and not lib.supports(arm):
                    continue
                best = None
                for _ in range(3):
                    secs, tb_out, st, agg = lib.pusch_concurrent(arm, pool if arm.startswith("cuda") else "sw", llrs, tbs // 8,
                                                                 bg, Qm, layers, aggregate_blocks=max(blocks, 2), warm_reps=3)
                    assert (st[:, :, 0, 0] == 1).all() and (tb_out == tb).all(), (arm, K)
                    best = secs if best is None else min(best, secs)
                key = arm + ("-%dblocks" % blocks if arm == "cuda-aggregated" else "")
                case["arms"].setdefault(key, {})["K=%d" % K] = {
                    "tb_per_s": K * reps / best, "gbps": K * reps * tbs / best / 1e9, "mean_tb_latency_us": best / reps * 1e6,
                    "tbs_per_launch": (agg[1] / agg[0]) if agg[0] else None,
                    "us_per_tb": {"join_wait": agg[3] / agg[1] / 1e3, "stage": agg[4] / agg[1] / 1e3,
                                  "follow": agg[5] / agg[1] / 1e3, "drain": agg[6] / agg[1] / 1e3,
                                  "launch_us_per_launch": agg[7] / agg[0] / 1e3} if agg[0] else None}
        out["cases"].append(case)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
