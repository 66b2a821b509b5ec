"""GPU parity of the soft demodulator + descrambler (ldpc_cuda_demodulate_soft, ldpc_cuda_descramble,
ldpc_cuda_demodulate_batch_device) against the oracle restatement of demodulation_mapper_impl::demodulate_soft
(demodulation_mapper_impl.cpp:36-106, generic per-symbol code) and pusch_demodulator_impl::descramble
(pusch_demodulator_impl.cpp:100-152): bit-exact. Against the reference's AVX2 kernels (which differ from the reference's
own generic code): at most one LLR step on fewer than 1e-4 of the values. Floating point, tolerance stated here: 0 vs the
generic path."""
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


@pytest.mark.parametrize("mod_order,pi2", [(1, False), (1, True), (2, False), (4, False), (6, False), (8, False)])
def test_demodulate_soft_bit_exact(dec, mod_order, pi2):
    rng = np.random.default_rng(200 + mod_order)
    for noise_var, n in [(0.3, 30001), (0.03, 8192), (0.003, 77)]:
        sym, nv = O.random_symbols(rng, mod_order, n, noise_var)
        got = dec.demodulate_soft(mod_order, sym, nv, pi2)
        want = O.port().demodulate_soft(mod_order, sym, nv, pi2)
        assert (got == want).all(), np.flatnonzero(got != want)[:8]
        ref = O.ref()
        if ref is not None:
            simd = ref.demodulate_soft(mod_order, sym, nv, pi2, 0)
            diff = np.abs(simd.astype(np.int16) - got)
            assert diff.max() <= 1 and (diff != 0).mean() < 1e-3


def _tie_inputs(rng, count=64):
    """QPSK axis values x (noise variance 1) whose quantiser input RN(RN(l / 24) * 120), l = RN(2 sqrt2 x), lands exactly on
    k + 0.5: std::round goes away from zero there, round-to-nearest-even does not."""
    c = np.float32(2.0) * np.float32(1.41421356237309504880)
    found = []
    while sum(f.size for f in found) < count:
        x = rng.uniform(-9.0, 9.0, 4_000_000).astype(np.float32)
        b = (c * x / np.float32(24.0)) * np.float32(120.0)
        assert b.dtype == np.float32
        found.append(x[(np.abs(b) < 120) & (np.abs(b) * 2 % 2 == 1)])
    return np.concatenate(found)[:count]


@pytest.mark.parametrize("mod_order", [1, 2, 4, 6, 8])
def test_demodulate_soft_special_values(dec, mod_order):
    """Rounding ties, signed zeros, subnormals, far outliers, infinities and NaN in the symbols; zero, negative,
    subnormal, huge, infinite and NaN noise variances (log_likelihood_ratio.cpp:87-96 clip + round, the nv <= 0 branches
    of demodulation_mapper_*.cpp). Bit-exact with the oracle port on the same platform."""
    rng = np.random.default_rng(900 + mod_order)
    ties = _tie_inputs(rng)
    vals = np.concatenate([ties, -ties, np.float32([0.0, -0.0, 1e-45, -1e-45, 1e-39, -3e-39, 1e-30, 0.5, -0.5, 1.0, 3.25,
                                                    -77.0, 1e4, -1e4, 1e7, -1e7, np.inf, -np.inf, np.nan]),
                           (rng.integers(-16, 17, 200) / np.sqrt(O.MOD_NORM[mod_order])).astype(np.float32),  # boundaries
                           rng.normal(0, 1, 300).astype(np.float32)])
    nvs = np.float32([1.0, 0.0, -0.0, -1.0, 1e-45, 1e-39, 1e-20, 0.01, 3.0, 1e20, 3e38, np.inf, np.nan])
    re = np.repeat(vals, nvs.size)
    nv = np.tile(nvs, vals.size)
    im = rng.permutation(re)
    for a, b in ((re, im), (im, re), (re, np.zeros_like(re))):
        sym = np.empty(a.size, np.complex64)
        sym.real, sym.imag = a, b
        with np.errstate(all="ignore"):
            want = O.port().demodulate_soft(mod_order, sym, nv)
        got = dec.demodulate_soft(mod_order, sym, nv)
        bad = np.flatnonzero(got != want)
        assert bad.size == 0, (bad[:8], got[bad[:8]], want[bad[:8]])


@pytest.mark.parametrize("c_init,n", [(0, 1), (1, 31), (0x7fffffff, 33), ((0x4601 << 15) + 935, 100001),
                                      (0x12345678, 8192), (77, 8193), ((0xffff << 15) + 1023, 1362816),
                                      (0x2468ace, (1 << 21) + 70001)])  # beyond the precomputed x1 words
def test_descramble_bit_exact(dec, c_init, n):
    rng = np.random.default_rng(n)
    llrs = rng.integers(-127, 128, n).astype(np.int8)
    got = dec.descramble(c_init, llrs)
    assert (got == O.port().descramble(c_init, llrs)).all()
    ref = O.ref()
    if ref is not None:
        assert (got == ref.descramble(c_init, llrs)).all()


def test_demodulate_batch_device_heterogeneous_jobs(dec):
    """Several codewords of different modulations in one launch, unaligned output offsets, descrambling on and off."""
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(3)
    cases = [(8, False, 1, 5000, 0x11111), (6, False, 1, 4099, 0x7abcdef), (4, False, 0, 2500, 0), (2, False, 1, 9001, 5),
             (1, True, 1, 8193, 0x55555), (1, False, 1, 100, 1), (0, False, 1, 20000, 0x3333)]
    jobs = np.zeros(len(cases), pkg.DEMOD_DESC)
    syms, nvs, want, so, lo = [], [], [], 0, 3
    in_llrs = rng.integers(-127, 128, 20000).astype(np.int8)
    for i, (q, pi2, scr, n, c_init) in enumerate(cases):
        j = jobs[i]
        j["mod_order"], j["pi2_bpsk"], j["descramble"], j["nof_symbols"], j["c_init"] = q, int(pi2), scr, n, c_init
        j["llr_offset"] = lo
        if q:
            sym, nv = O.random_symbols(rng, q, n, 0.05)
            j["symbol_offset"] = so
            syms.append(sym)
            nvs.append(nv)
            so += n
            w = O.port().demodulate_soft(q, sym, nv, pi2)
        else:
            j["symbol_offset"] = 0
            w = in_llrs
        want.append((lo, O.port().descramble(c_init, w) if scr else w))
        lo += w.size + (i % 3)
    sym_d = torch.from_numpy(np.concatenate(syms).view(np.float32)).to(dev)
    nv_d = torch.from_numpy(np.concatenate(nvs)).to(dev)
    in_d = torch.from_numpy(in_llrs).to(dev)
    out_d = torch.full((lo + 8,), 0x55, dtype=torch.int8, device=dev)
    torch.cuda.synchronize()
    dec.demodulate_batch_device(jobs, sym_d.data_ptr(), nv_d.data_ptr(), in_d.data_ptr(), out_d.data_ptr())
    dec.synchronize()
    out = out_d.cpu().numpy()
    end = 0
    for off, w in want:
        assert (out[end:off] == 0x55).all()  # gaps between the jobs untouched
        assert (out[off:off + w.size] == w).all(), off
        end = off + w.size
    assert (out[end:] == 0x55).all()


@pytest.mark.parametrize("nof_jobs", [40, 80, 160])
def test_demodulate_batch_device_many_jobs(dec, nof_jobs):
    """Launches large enough for 2, 4 and 8 spans of 8192 LLRs per block (the scrambler's jump-ahead is then paid once per
    warp and 2048 / 4096 / 8192 chips): ragged lengths, every modulation, a different c_init per job."""
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(40 + nof_jobs)
    nsym = 110003
    base = {}
    for q in (1, 2, 4, 6, 8):
        sym, nv = O.random_symbols(rng, q, nsym, 0.02)
        base[q] = (sym, nv, O.port().demodulate_soft(q, sym, nv))
    in_llrs = rng.integers(-127, 128, nsym).astype(np.int8)
    order = [8, 8, 8, 6, 8, 4, 8, 2, 8, 1, 8, 0]
    sym_all = np.concatenate([base[q][0] for q in (1, 2, 4, 6, 8)])
    nv_all = np.concatenate([base[q][1] for q in (1, 2, 4, 6, 8)])
    sym_off = {q: i * nsym for i, q in enumerate((1, 2, 4, 6, 8))}
    jobs = np.zeros(nof_jobs, pkg.DEMOD_DESC)
    want, lo = [], 0
    for i in range(nof_jobs):
        q = order[i % len(order)]
        n = nsym - 97 * (i % 5)  # ragged: the last block of a job ends inside a span
        j = jobs[i]
        j["mod_order"], j["descramble"], j["nof_symbols"], j["c_init"] = q, 1, n, (0x1357 * (i + 1)) & 0x7fffffff
        j["symbol_offset"] = sym_off[q] if q else 0
        j["llr_offset"] = lo
        w = base[q][2][:n * q] if q else in_llrs[:n]
        want.append((lo, O.port().descramble(int(j["c_init"]), w)))
        lo += (w.size + 15) // 8 * 8
    sym_d = torch.from_numpy(sym_all.view(np.float32)).to(dev)
    nv_d = torch.from_numpy(nv_all).to(dev)
    in_d = torch.from_numpy(in_llrs).to(dev)
    out_d = torch.full((lo,), 0x55, dtype=torch.int8, device=dev)
    torch.cuda.synchronize()
    dec.demodulate_batch_device(jobs, sym_d.data_ptr(), nv_d.data_ptr(), in_d.data_ptr(), out_d.data_ptr())
    dec.synchronize()
    out = out_d.cpu().numpy()
    for i, (off, w) in enumerate(want):
        bad = np.flatnonzero(out[off:off + w.size] != w)
        assert bad.size == 0, (i, int(jobs[i]["mod_order"]), bad[:8])
        assert (out[off + w.size:(off + w.size + 15) // 8 * 8] == 0x55).all()


def test_descramble_with_uci_placeholders(dec):
    """pusch_demodulator_impl::descramble with config.placeholders (pusch_demodulator_impl.cpp:100-152): codewords are
    demodulated + descrambled as plain data in one launch, ldpc_cuda_descramble_placeholders_device then amends the
    placeholder bits in HBM. Against the oracle restatement (pinned to the reference's own function on the CPU) and, when
    oracle/_ref is built, against that function directly."""
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(77)
    cases = [(2, 1, 600, [0, 5, 6, 100, 599]), (4, 2, 400, [3, 4, 5, 200, 399]), (6, 1, 1366, [1, 7, 682, 1365]),
             (8, 4, 300, [0, 1, 2, 150, 299]), (8, 1, 3000, [1024, 1025, 2047, 2048]), (1, 1, 500, [0, 2, 4, 250, 498]),
             (6, 3, 911, [0, 455, 910]), (4, 1, 50, []), (8, 2, 9000, list(range(0, 9000, 7)))]
    jobs = np.zeros(len(cases), pkg.DEMOD_DESC)
    ph = np.zeros(len(cases), pkg.PLACEHOLDER_DESC)
    syms, nvs, res_all, want, so, lo = [], [], [], [], 0, 0
    for i, (q, layers, nof_re, res) in enumerate(cases):
        n = nof_re * layers
        rnti, n_id = int(rng.integers(1, 65536)), int(rng.integers(0, 1024))
        sym, nv = O.random_symbols(rng, q, n, 0.05)
        j = jobs[i]
        j["mod_order"], j["descramble"], j["nof_symbols"], j["c_init"] = q, 1, n, rnti * 32768 + n_id
        j["symbol_offset"], j["llr_offset"] = so, lo
        p = ph[i]
        p["llr_offset"], p["c_init"], p["nof_llrs"], p["first_re"] = lo, rnti * 32768 + n_id, n * q, len(res_all)
        p["nof_re"], p["mod_order"], p["nof_layers"] = len(res), q, layers
        res_all += res
        llrs = O.port().demodulate_soft(q, sym, nv)
        w = O.port().descramble_placeholders(rnti * 32768 + n_id, llrs, q, layers, res)
        if O.ref() is not None:
            assert (w == O.ref().descramble_placeholders(rnti, n_id, llrs, q, layers, res)).all()
        want.append((lo, w))
        syms.append(sym)
        nvs.append(nv)
        so += n
        lo += n * q + 5
    sym_d = torch.from_numpy(np.concatenate(syms).view(np.float32)).to(dev)
    nv_d = torch.from_numpy(np.concatenate(nvs)).to(dev)
    out_d = torch.full((lo,), 0x55, dtype=torch.int8, device=dev)
    torch.cuda.synchronize()
    dec.demodulate_batch_device(jobs, sym_d.data_ptr(), nv_d.data_ptr(), 0, out_d.data_ptr())
    dec.descramble_placeholders_device(ph, np.array(res_all, np.uint16), out_d.data_ptr())
    dec.synchronize()
    out = out_d.cpu().numpy()
    for i, (off, w) in enumerate(want):
        bad = np.flatnonzero(out[off:off + w.size] != w)
        assert bad.size == 0, (i, cases[i][:3], bad[:8])
        assert (out[off + w.size:off + w.size + 5] == 0x55).all()
    # the host-buffer entry (ldpc_cuda_descramble_uci) on the first cases
    for i, (q, layers, nof_re, res) in enumerate(cases[:4]):
        c_init = int(jobs[i]["c_init"])
        llrs = rng.integers(-127, 128, q * layers * nof_re).astype(np.int8)
        assert (dec.descramble_uci(c_init, llrs, q, layers, res) ==
                O.port().descramble_placeholders(c_init, llrs, q, layers, res)).all(), i
    # argument checks: RE list shorter than a descriptor claims, unsupported modulation order
    bad = ph[:1].copy()
    bad["first_re"] = len(res_all)
    with pytest.raises(pkg.LdpcCudaError):
        dec.descramble_placeholders_device(bad, np.array(res_all, np.uint16), out_d.data_ptr())
    bad = ph[:1].copy()
    bad["mod_order"] = 3
    with pytest.raises(pkg.LdpcCudaError):
        dec.descramble_placeholders_device(bad, np.array(res_all, np.uint16), out_d.data_ptr())


QAM_BITS = {2: 1, 4: 2, 6: 3, 8: 4}


def _modulate(bits, q):
    """TS 38.211 5.1.3-5.1.6 mapping (bit 2k -> real axis, bit 2k+1 -> imaginary axis, bit 0/1 = sign)."""
    b = bits.reshape(-1, q).astype(np.int64)
    nb = QAM_BITS[q]

    def axis(cols):
        v = 2 - (1 - 2 * cols[:, nb - 1]) if nb > 1 else np.ones(len(cols), np.int64)
        step = 2
        for k in range(nb - 2, 0, -1):
            step *= 2
            v = step - (1 - 2 * cols[:, k]) * v
        return (1 - 2 * cols[:, 0]) * v
    re, im = axis(b[:, 0::2]), axis(b[:, 1::2])
    return ((re + 1j * im) / np.sqrt(O.MOD_NORM[q])).astype(np.complex64)


@pytest.mark.parametrize("tbs,Qm,layers,nsym,esn0", [(100008, 6, 2, 30000, 22.0), (25000 // 8 * 8 + 8, 4, 1, 9000, 14.0),
                                                    (100008, 8, 2, 17000, 30.0)])
def test_symbols_to_transport_block_without_leaving_the_device(dec, tbs, Qm, layers, nsym, esn0):
    """Whole receive chain on the device: equalised symbols -> demodulate + descramble -> rate dematch + decode + CRC ->
    transport block assembly + CRC24A. The transmitter (segmentation, encoder, rate matcher, scrambler, QAM mapper) is the
    oracle / numpy. The assembled block equals the transmitted one."""
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(tbs % 1000)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    try:
        tb, cws, meta = O.build_transport_block(rng, tbs, 1, Qm, layers, nsym)
    except AssertionError:
        pytest.skip("not a 3GPP transport block size")
    descs, crc_bits = pkg.segment(tbs, 1, Qm, layers, nsym)
    C = descs.size
    tx = np.concatenate([O.port().rate_match(cws[r], int(meta[r, 2]), 0, Qm, 0, int(meta[r, 3])) for r in range(C)])
    c_init = (0x1234 << 15) + 77
    chips = O.port().gold_sequence(c_init, tx.size)
    sym = _modulate(tx ^ chips, Qm)
    sigma2 = 10.0 ** (-esn0 / 10.0)
    sym = (sym + (rng.normal(0, 1, sym.size) + 1j * rng.normal(0, 1, sym.size)) * np.sqrt(sigma2 / 2)).astype(np.complex64)
    nv = np.full(sym.size, sigma2, np.float32)
    ro = so = oo = 0
    for r in range(C):
        d = descs[r]
        Z = int(d["lifting_size"])
        d["early_stop"], d["new_data"], d["in_offset"], d["soft_offset"], d["out_offset"] = 1, 1, ro, so, oo
        d["in_length"] = 66 * Z
        ro += int(d["rm_length"])
        so += 66 * 384
        oo += (22 * Z + 7) // 8
    assert ro == tx.size
    dec.softbuffer_reserve(so)
    job = np.zeros(1, pkg.DEMOD_DESC)
    job["mod_order"], job["descramble"], job["nof_symbols"], job["c_init"] = Qm, 1, sym.size, c_init
    tbd = np.zeros(1, pkg.TB_DESC)
    tbd["nof_codeblocks"], tbd["tbs_bits"], tbd["nof_cb_crc_bits"] = C, tbs, int(crc_bits[0])
    sym_d = torch.from_numpy(sym.view(np.float32)).to(dev)
    nv_d = torch.from_numpy(nv).to(dev)
    llr_d = torch.zeros(tx.size, dtype=torch.int8, device=dev)
    out_d = torch.zeros(oo, dtype=torch.uint8, device=dev)
    res_d = torch.zeros(C, dtype=torch.int32, device=dev)
    tb_d = torch.zeros(tbs // 8 + 3, dtype=torch.uint8, device=dev)
    tbres_d = torch.zeros(2, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    dec.demodulate_batch_device(job, sym_d.data_ptr(), nv_d.data_ptr(), 0, llr_d.data_ptr())
    dec.dematch_decode_batch_device(descs, llr_d.data_ptr(), out_d.data_ptr(), res_d.data_ptr())
    dec.tb_assemble_device(tbd, descs, out_d.data_ptr(), res_d.data_ptr(), tb_d.data_ptr(), tbres_d.data_ptr())
    dec.synchronize()
    llr = llr_d.cpu().numpy()
    assert (llr == O.port().descramble(c_init, O.port().demodulate_soft(Qm, sym, nv))).all()
    tbres = tbres_d.cpu().numpy().view(pkg.TB_RESULT)
    assert int(tbres[0]["tb_crc_ok"]) == 1 and int(tbres[0]["nof_cb_failed"]) == 0
    assert (tb_d.cpu().numpy()[:tbs // 8] == tb).all()
