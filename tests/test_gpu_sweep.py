"""Seeded randomized property sweep through the FUSED entry (ldpc_cuda_dematch_decode_batch: rate dematch + HARQ combine in
the arena -> layered min-sum -> CRC early stop) over ALL 102 (base graph, lifting size) combinations of TS 38.212 5.3.2
(SURVEY.md 8(d) cfg 2 / P1), against the plain-C oracle (pusch_decoder_impl.cpp:156-201 per codeblock).

Every case draws E, rv, modulation order, Nref (limited-buffer rate matching), filler bits, CRC polynomial, early stop,
new data / retransmission and skip_decode at random; the arena holds arbitrary prior contents; every call mixes all base
graphs and lifting sizes, so the two codeblocks that share a thread pair in the kernel differ in everything except
(BG, Z). The iteration limit takes the values {1, 6, 10}. A second call retransmits a third of the codeblocks on top of
the soft bits the first call left in HBM. Soft buffers, message bits, crc_ok and iteration counts must equal the oracle's.
A failure prints the seed and the case."""
import os
import time

import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()
pytestmark = pytest.mark.gpu

SEED = int(os.environ.get("LDPC_SWEEP_SEED", "20231"))
CASES_PER_SHAPE = int(os.environ.get("LDPC_SWEEP_CASES", "10"))  # x 102 shapes x 3 iteration limits = 3060 first transmissions
ALL_Z = sorted(a * (1 << j) for a in (2, 3, 5, 7, 9, 11, 13, 15) for j in range(8) if a * (1 << j) <= 384)


def _draw(rng, bg, Z):
    """One codeblock: parameters, the rate-matched LLRs of its first transmission and the transmitter state."""
    K, Ns = O.BG_K[bg], O.BG_N_SHORT[bg]
    N = Ns * Z
    fits = [p for p, nb in ((O.CRC16, 16), (O.CRC24A, 24), (O.CRC24B, 24)) if nb + 8 <= K * Z]
    crc = int(rng.choice(fits)) if fits else O.CRC16  # BG2 Z=2 carries 20 bits
    crc_bits = 16 if crc == O.CRC16 else 24
    max_f = min((K - 2) * Z - 1, K * Z - crc_bits - 1)
    F = int(rng.integers(0, max(1, min(max_f, 4 * Z)))) if rng.integers(0, 3) else 0
    Qm = int(rng.choice([1, 2, 4, 6, 8]))
    Nref = 0 if rng.integers(0, 3) else min(25344, int(rng.integers((K + 4) * Z, N + Z)))  # some beyond N: no limit then
    kind = int(rng.integers(0, 10))  # 0: noise only, 1: empty transmission, 2: saturated, else a codeword in noise
    lo, hi = int(1.05 * K * Z), int(1.6 * N)
    E = int(rng.integers(lo, hi)) // Qm * Qm
    if kind == 1:
        E = 0
    msg, cw = O.make_codeword(rng, bg, Z, F, crc)
    c = dict(bg=bg, Z=Z, N=N, K=K, F=F, crc=crc, Qm=Qm, Nref=Nref, cw=cw, kind=kind, early=int(rng.integers(0, 4) != 0),
             esn0=float(rng.uniform(-3.0, 5.0)) - (2.0 if bg == 2 else 0.0), kb=(K * Z + 7) // 8)
    c["tx"] = [_transmission(rng, c, E, int(rng.integers(0, 4)) if rng.integers(0, 3) == 0 else 0)]
    return c


def _transmission(rng, c, E, rv):
    if c["kind"] == 0:
        rm = rng.integers(-120, 121, E).astype(np.int8)
    else:
        rm = O.awgn_llrs(rng, O.port().rate_match(c["cw"], E, rv, c["Qm"], c["Nref"], c["F"]), c["esn0"],
                         range_limit=6.0 if c["kind"] == 2 else 20.0)
    return dict(E=E, rv=rv, rm=rm)


def _run_call(dec, cbs, t, max_iters):
    """One fused call over cbs (transmission t of each); checks everything against the oracle."""
    n = len(cbs)
    descs = np.zeros(n, pkg.CB_DESC)
    ro = 0
    for i, c in enumerate(cbs):
        x, d = c["tx"][t], descs[i]
        d["base_graph"], d["lifting_size"], d["nof_filler_bits"], d["crc_poly"] = c["bg"], c["Z"], c["F"], c["crc"]
        d["early_stop"], d["rv"], d["mod_order"], d["Nref"], d["new_data"] = c["early"], x["rv"], c["Qm"], c["Nref"], c["new"][t]
        d["rm_length"], d["in_offset"], d["out_offset"], d["soft_offset"] = x["E"], ro, c["oo"], c["so"]
        d["in_length"], d["skip_decode"] = c["N"], c["skip"][t]
        x["ro"] = ro
        ro += x["E"] + int(i % 3 == 0)  # odd starts now and then
    rm_all = np.zeros(ro + 16, np.int8)
    for c in cbs:
        x = c["tx"][t]
        rm_all[x["ro"]:x["ro"] + x["E"]] = x["rm"]
    out = np.full(cbs[-1]["oo"] + cbs[-1]["kb"], 0xa5, np.uint8)
    out, res = dec.dematch_decode_batch(descs, rm_all, out)
    arena = dec.softbuffer_read(0, cbs[-1]["so"] + cbs[-1]["N"])
    for i, c in enumerate(cbs):
        x = c["tx"][t]
        tag = "LDPC_SWEEP_SEED=%d max_iters=%d tx=%d cb=%d %s" % (
            SEED, max_iters, t, i, {k: c[k] for k in ("bg", "Z", "F", "crc", "Qm", "Nref", "kind", "early", "new", "skip")})
        O.port().rate_dematch(c["soft"], x["rm"], c["new"][t], x["rv"], c["Qm"], c["Nref"], c["F"])
        got = arena[c["so"]:c["so"] + c["N"]]
        assert (got == c["soft"]).all(), (tag, x["E"], x["rv"], np.nonzero(got != c["soft"])[0][:8])
        o = out[c["oo"]:c["oo"] + c["kb"]]
        if c["skip"][t]:
            assert (o == 0xa5).all(), tag
            continue
        soft = c["soft"]
        if c["early"]:
            bits, it = O.port().decode(c["bg"], c["Z"], soft, fillers=c["F"], crc=c["crc"], max_iters=max_iters, out_init=0xa5)
            ok, iters = it > 0, (it if it else max_iters)
        else:  # decode_cblk without early stop: all iterations, one CRC check at the end
            bits, _ = O.port().decode(c["bg"], c["Z"], soft, fillers=c["F"], crc=O.CRC_NONE, max_iters=max_iters, out_init=0xa5)
            ok, iters = O.port().crc(c["crc"], bits, c["K"] * c["Z"] - c["F"]) == 0, max_iters
        if not soft.any():  # nothing received yet: the reference does not decode (ldpc_decoder_impl.cpp:66-71)
            assert not res["crc_ok"][i], tag
            continue
        if bool(res["crc_ok"][i]) != ok or int(res["iterations"][i]) != iters:
            raise AssertionError("%s: crc_ok %d iterations %d, oracle %d %d (E %d rv %d)" % (
                tag, res["crc_ok"][i], res["iterations"][i], ok, iters, x["E"], x["rv"]))
        assert (o == bits).all(), tag
    return n


@pytest.mark.parametrize("flags,per_shape", [(0, CASES_PER_SHAPE), (pkg.FLAG_FORCE_SCALAR, 2)], ids=["auto", "scalar"])
def test_fused_entry_all_shapes_randomized(flags, per_shape):
    dec = pkg.LdpcCuda(device=0, max_codeblocks=4096, flags=flags)
    try:
        _sweep(dec, per_shape)
        if not flags & pkg.FLAG_FORCE_SCALAR:
            # every (BG, Z) of the sweep ran on the fused tensor-memory kernel, none fell back to dematch + general decoder
            assert dec.general_kernel_launch_count == 0
    finally:
        dec.close()


def _sweep(dec, per_shape):
    t0 = time.time()
    rng = np.random.default_rng(SEED + per_shape)
    total = 0
    shapes = [(bg, Z) for bg in (1, 2) for Z in ALL_Z]
    assert len(shapes) == 102
    for max_iters in (1, 6, 10):
        dec.set_algorithm(max_iters, 0.8, O.SCALE_GENERIC)
        cbs = []
        for bg, Z in shapes:
            for _ in range(per_shape):
                cbs.append(_draw(rng, bg, Z))
        order = rng.permutation(len(cbs))  # neighbours in the call (and so in the arena) are unrelated shapes
        cbs = [cbs[k] for k in order]
        so = oo = 0
        for i, c in enumerate(cbs):
            so += int(rng.integers(0, 3)) * int(rng.integers(1, 16))  # unaligned soft buffers too
            c["so"], c["oo"] = so, oo
            so += c["N"]
            oo += c["kb"]
            # first transmission: new data, or a "retransmission" on top of arbitrary prior contents, some skipped
            c["new"] = [int(rng.integers(0, 4) != 0)]
            c["skip"] = [int(c["new"][0] == 0 and rng.integers(0, 3) == 0)]
            c["soft"] = rng.integers(-120, 121, c["N"]).astype(np.int8)
            c["soft"][(c["K"] - 2) * c["Z"] - c["F"]:(c["K"] - 2) * c["Z"]] = 127
        dec.softbuffer_reserve(so + 64)
        prior = np.zeros(so, np.int8)
        for c in cbs:
            prior[c["so"]:c["so"] + c["N"]] = c["soft"]
        dec.softbuffer_write(0, prior)
        total += _run_call(dec, cbs, 0, max_iters)
        # second call: a third of the codeblocks is retransmitted with another redundancy version (and some of those only
        # combined, skip_decode), on top of what the first call left in HBM
        again = [c for c in cbs if rng.integers(0, 3) == 0]
        for c in again:
            E = int(rng.integers(c["Z"], int(1.2 * c["N"]))) // c["Qm"] * c["Qm"]
            c["tx"].append(_transmission(rng, c, E, int(rng.integers(0, 4))))
            c["new"].append(0)
            c["skip"].append(int(rng.integers(0, 4) == 0))
        last = max(again, key=lambda c: c["so"])
        again = [c for c in again if c is not last] + [last]  # _run_call sizes its buffers from the last entry
        total += _run_call(dec, again, 1, max_iters)
    elapsed = time.time() - t0
    print("sweep: %d codeblock cases in %.1f s (seed %d)" % (total, elapsed, SEED))
    assert total >= 200 * per_shape
