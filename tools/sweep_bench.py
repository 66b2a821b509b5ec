#!/usr/bin/env python3
"""Fixed-6-iteration decode throughput over base graphs and lifting sizes (device-resident random LLRs, no CRC), plus the
latency of the per-codeblock drop-in call. Development / profiling helper: writes a markdown table.

    python tools/sweep_bench.py > profiles/r01_sweep_bg_z.md
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def main():
    import torch
    pkg = ge.load_package()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    sf = float(os.environ.get("LDPC_SWEEP_SF", "0.8"))  # != 0.8: the table-driven scaling of the fast kernel
    dec = pkg.LdpcCuda(device=0, max_codeblocks=1 << 17, max_iterations=6, scaling_factor=sf)
    stream = torch.cuda.ExternalStream(dec.stream, device=dev)
    print("| BG | Z | codeblocks | info bits / CB | ms | decoded info Gb/s | EMU/s (1e12) |")
    print("|---|---|---|---|---|---|---|")
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    for bg in (1, 2):
        for Z in (384, 352, 320, 256, 208, 192, 128, 96, 64, 52, 32, 20, 8, 2):
            K, N = (22 if bg == 1 else 10) * Z, (66 if bg == 1 else 50) * Z
            edges = 316 if bg == 1 else 197
            n = int(min(1 << 17, max(4096, (1 << 29) // N)))
            llr = torch.randint(-40, 41, (n, N), generator=g, device=dev, dtype=torch.int8)
            out = torch.zeros((n, (K + 7) // 8), dtype=torch.uint8, device=dev)
            res = torch.zeros(n, dtype=torch.int32, device=dev)
            descs = pkg.make_descs(n, bg, Z, crc_poly=pkg.CRC_NONE)
            torch.cuda.synchronize()
            for _ in range(2):
                dec.decode_batch_device(descs, llr.data_ptr(), out.data_ptr(), res.data_ptr())
            dec.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            reps = 3
            for _ in range(reps):
                dec.decode_batch_device(descs, llr.data_ptr(), out.data_ptr(), res.data_ptr())
            e1.record(stream)
            dec.synchronize()
            ms = e0.elapsed_time(e1) / reps
            print("| %d | %d | %d | %d | %.3f | %.2f | %.3f |" % (bg, Z, n, K, ms, n * K / ms / 1e6,
                                                               n * 6.0 * edges * Z / ms / 1e9))
            del llr, out, res
    # per-codeblock drop-in latency (ldpc_cuda_decode: H2D 25 KB + one kernel + D2H 1 KB, synchronous)
    print()
    print("| drop-in call | mean latency (us) |")
    print("|---|---|")
    rng = np.random.default_rng(0)
    for bg, Z, crc, what in ((1, 384, pkg.CRC_NONE, "BG1 Z=384, 6 iterations, no CRC"),
                             (2, 384, pkg.CRC_NONE, "BG2 Z=384, 6 iterations, no CRC"),
                             (1, 52, pkg.CRC_NONE, "BG1 Z=52, 6 iterations, no CRC")):
        N = (66 if bg == 1 else 50) * Z
        x = rng.integers(-40, 41, N).astype(np.int8)
        for _ in range(20):
            dec.decode(bg, Z, x, crc=crc)
        t0 = time.perf_counter()
        reps = 200
        for _ in range(reps):
            dec.decode(bg, Z, x, crc=crc)
        print("| ldpc_cuda_decode, %s | %.1f |" % (what, (time.perf_counter() - t0) / reps * 1e6))
    dec.close()


if __name__ == "__main__":
    main()
