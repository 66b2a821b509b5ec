#!/usr/bin/env python3
"""Generates tests/golden/ldpc_golden.npz FROM THE COMPILED REFERENCE (oracle/_ref/libsrsran_ref.so, built from the sources
in /root/reference by oracle/Makefile). The reference's own MATLAB test-vector tarballs are not in its tree (SURVEY.md
8c), so these vectors - inputs plus what the reference's generic decoder / rate dematcher / CRC return for them - are the
committed known answers. Run in the build container only (needs /root/reference); the tests read the .npz.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracles as O  # noqa: E402

# (bg, Z, fillers, crc, nodes or 0 = full, esn0 dB, max_iters)
DECODE_CASES = [
    (1, 384, 0, O.CRC24B, 0, -1.0, 6), (1, 384, 0, O.CRC_NONE, 0, -1.0, 6), (1, 384, 16, O.CRC24B, 26, 7.0, 6),
    (1, 384, 0, O.CRC24A, 0, -2.6, 10), (1, 208, 28, O.CRC24B, 40, 0.5, 6), (1, 36, 0, O.CRC16, 0, -1.5, 6),
    (1, 2, 0, O.CRC_NONE, 0, 2.0, 2), (2, 384, 0, O.CRC24B, 0, -4.0, 6), (2, 240, 12, O.CRC16, 30, -1.0, 6),
    (2, 15, 0, O.CRC16, 0, -3.0, 6), (2, 7, 0, O.CRC_NONE, 12, 1.0, 1), (2, 52, 5, O.CRC24A, 0, -5.5, 6),
]
# (bg, Z, E, rv, Qm, Nref, F, retransmissions)
DEMATCH_CASES = [
    (1, 16, 300, 0, 2, 0, 12, 2), (1, 16, 2000, 0, 2, 700, 12, 1), (1, 16, 300, 2, 2, 0, 12, 1), (2, 12, 704, 2, 8, 0, 0, 2),
    (1, 384, 8960, 0, 8, 0, 16, 3), (1, 384, 8992, 3, 8, 0, 16, 1), (1, 384, 30000, 1, 2, 16896, 40, 1),
    (2, 384, 6000, 0, 6, 0, 0, 3), (1, 56, 1400, 3, 4, 2000, 12, 2), (2, 9, 200, 1, 1, 0, 0, 2),
]


def main():
    ref = O.ref()
    assert ref is not None, "oracle/_ref/libsrsran_ref.so is missing: run `make -C oracle ref` in the build container"
    rng = np.random.default_rng(20230501)
    out = {}
    for i, (bg, Z, F, crc, nodes, esn0, iters) in enumerate(DECODE_CASES):
        msg, cw = O.make_codeword(rng, bg, Z, F, crc if crc else O.CRC24B)
        llr = O.awgn_llrs(rng, cw, esn0)
        if nodes:
            llr = llr[:nodes * Z].copy()
        bits, it = ref.decode(bg, Z, llr, fillers=F, crc=crc, max_iters=iters, kind="generic")
        out["dec%d_llr" % i] = llr
        out["dec%d_bits" % i] = bits
        out["dec%d_meta" % i] = np.array([bg, Z, F, crc, iters, it], np.int32)
        if ref.supports("avx2"):
            abits, ait = ref.decode(bg, Z, llr, fillers=F, crc=crc, max_iters=iters, kind="avx2")
            out["dec%d_avx2_bits" % i] = abits
            out["dec%d_avx2_it" % i] = np.array([ait], np.int32)
    for i, (bg, Z, E, rv, Qm, Nref, F, retx) in enumerate(DEMATCH_CASES):
        N = O.BG_N_SHORT[bg] * Z
        soft = rng.integers(-120, 121, N).astype(np.int8) if i % 2 else np.zeros(N, np.int8)
        out["rm%d_prior" % i] = soft.copy()
        out["rm%d_meta" % i] = np.array([bg, Z, E, rv, Qm, Nref, F, retx], np.int32)
        for t in range(1 + retx):
            x = rng.integers(-120, 121, E).astype(np.int8)
            ref.rate_dematch(soft, x, t == 0, bg, Z, rv, Qm, Nref, F, kind="generic")
            out["rm%d_in%d" % (i, t)] = x
            out["rm%d_out%d" % (i, t)] = soft.copy()
    # CRC known answers over ragged bit lengths
    data = rng.integers(0, 256, 1100).astype(np.uint8)
    lens = [1, 7, 8, 9, 24, 25, 100, 1055, 8424, 8448, 8800]
    out["crc_data"] = data
    out["crc_lens"] = np.array(lens, np.int32)
    out["crc_vals"] = np.array([[ref.crc(p, data, n) for n in lens] for p in (O.CRC16, O.CRC24A, O.CRC24B)], np.uint32)
    path = os.path.join(HERE, "ldpc_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
