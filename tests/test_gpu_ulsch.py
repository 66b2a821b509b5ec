"""GPU parity of the UL-SCH demultiplexer (ldpc_cuda_ulsch_demultiplex*, TS 38.212 6.2.7) against the oracle walk, which is
pinned against the reference's ulsch_demultiplex_impl (tests/test_ulsch_cpu.py): every soft bit of SCH data, HARQ-ACK, CSI
part 1 and CSI part 2, over the reference's message-information space (HARQ-ACK 0 / 1 / 2 / more bits with and without
reserved elements, CSI part 1 / 2, every modulation, 1-4 layers), through the host entry and through the batched device
entry; and the chain demodulate -> descramble -> demultiplex -> rate dematch + decode without leaving the device."""
import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dec():
    d = pkg.LdpcCuda(device=0, max_codeblocks=256)
    yield d
    d.close()


def _desc(cfg, G_ack, G_c1, G_c2):
    return pkg.ulsch_desc(cfg["mod_order"], cfg["nof_layers"], cfg["nof_prb"], cfg["start_symbol"], cfg["nof_symbols"],
                          cfg["dmrs_type"], cfg["cdm"], cfg["dmrs_mask"], cfg["rvd"], G_ack, G_c1, G_c2)


def _valid_cases(rng, n):
    out = []
    for cfg, O_ack, G_ack, O_c1, G_c1, O_c2, G_c2 in O.ulsch_cases(rng, 4 * n):
        sizes = pkg.ulsch_sizes(_desc(cfg, G_ack, G_c1, G_c2))
        if sizes is not None:
            out.append((cfg, G_ack, G_c1, G_c2, sizes))
        if len(out) == n:
            break
    return out


def test_host_entry_equals_oracle(dec):
    rng = np.random.default_rng(21)
    kinds = set()
    for cfg, G_ack, G_c1, G_c2, (n_in, n_sch) in _valid_cases(rng, 120):
        llrs = rng.integers(-120, 121, n_in).astype(np.int8)
        llrs[llrs == 0] = 1  # zeros in the output then only come from punctured reserved elements
        want = O.port().ulsch_demultiplex(cfg, G_ack, G_c1, G_c2, llrs)
        assert want is not None
        got = dec.ulsch_demultiplex(_desc(cfg, G_ack, G_c1, G_c2), llrs)
        for name, a, b in zip(("sch", "ack", "csi1", "csi2"), got, want):
            assert a.size == b.size and (a == b).all(), (name, cfg)
        kinds.add((bool(G_ack), bool(cfg["rvd"]), bool(G_c1), bool(G_c2)))
    assert len(kinds) >= 8  # with / without each field, reserved and unreserved HARQ-ACK


def test_batched_device_entry_equals_oracle(dec):
    import torch
    rng = np.random.default_rng(22)
    cases = _valid_cases(rng, 40)
    descs = np.zeros(len(cases), pkg.ULSCH_DESC)
    ins, wants = [], []
    io = so = uo = 0
    for k, (cfg, G_ack, G_c1, G_c2, (n_in, n_sch)) in enumerate(cases):
        llrs = rng.integers(-120, 121, n_in).astype(np.int8)
        ins.append(llrs)
        wants.append(O.port().ulsch_demultiplex(cfg, G_ack, G_c1, G_c2, llrs))
        d = _desc(cfg, G_ack, G_c1, G_c2)
        d["in_offset"], d["sch_offset"] = io, so
        d["harq_ack_offset"], d["csi_part1_offset"], d["csi_part2_offset"] = uo, uo + G_ack, uo + G_ack + G_c1
        descs[k] = d[0]
        io += n_in + 3   # unaligned starts
        so += n_sch + 5
        uo += G_ack + G_c1 + G_c2 + 1
    in_h = np.zeros(io, np.int8)
    for d, x in zip(descs, ins):
        in_h[int(d["in_offset"]):int(d["in_offset"]) + x.size] = x
    in_d = torch.from_numpy(in_h).cuda()
    sch_d = torch.full((so,), 77, dtype=torch.int8, device="cuda")
    uci_d = torch.full((uo + 1,), 77, dtype=torch.int8, device="cuda")
    torch.cuda.synchronize()
    dec.ulsch_demultiplex_batch_device(descs, in_d.data_ptr(), sch_d.data_ptr(), uci_d.data_ptr())
    dec.synchronize()
    sch, uci = sch_d.cpu().numpy(), uci_d.cpu().numpy()
    for d, w in zip(descs, wants):
        s0 = int(d["sch_offset"])
        assert (sch[s0:s0 + w[0].size] == w[0]).all()
        assert (sch[s0 + w[0].size:s0 + w[0].size + 5] == 77).all()  # nothing written behind the field
        for off, part in ((int(d["harq_ack_offset"]), w[1]), (int(d["csi_part1_offset"]), w[2]),
                          (int(d["csi_part2_offset"]), w[3])):
            assert (uci[off:off + part.size] == part).all()


def _source_map(cfg, G_ack, G_c1, G_c2, n_in):
    """For every soft bit of the four demultiplexed fields, the codeword position it is read from (-1 = punctured element,
    zero-filled): the oracle walk run on the three byte planes of a position ramp."""
    idx = np.arange(1, n_in + 1, dtype=np.uint32)
    acc = None
    for p in range(3):
        plane = ((idx >> (8 * p)) & 0xff).astype(np.uint8).view(np.int8)
        out = O.port().ulsch_demultiplex(cfg, G_ack, G_c1, G_c2, plane)
        assert out is not None
        part = [x.view(np.uint8).astype(np.int64) << (8 * p) for x in out]
        acc = part if acc is None else [a + b for a, b in zip(acc, part)]
    return [a - 1 for a in acc]


QAM_BITS = {1: 1, 2: 1, 4: 2, 6: 3, 8: 4}


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


CHAIN = [
    # tbs, Qm, layers, prb, dmrs mask, O_ack, G_ack REs, rvd REs, G_csi1 REs, G_csi2 REs, Es/N0
    (100008, 6, 2, 106, 0b0000000000000100, 2, 60, 90, 150, 200, 24.0),   # reserved HARQ-ACK punctures SCH and CSI part 2
    (23976, 4, 1, 106, 0b0000100000000100, 4, 80, 0, 120, 0, 16.0),       # rate-matched HARQ-ACK, CSI part 1 only
    (23976, 2, 1, 273, 0b0000000000000100, 1, 40, 70, 0, 0, 10.0),        # one-bit HARQ-ACK: repetition placeholders
    (100008, 8, 2, 106, 0b0000000000000100, 0, 0, 120, 100, 300, 31.0),   # elements reserved, no HARQ-ACK sent
]


@pytest.mark.parametrize("tbs,Qm,layers,prb,mask,O_ack,re_ack,re_rvd,re_c1,re_c2,esn0", CHAIN)
def test_symbols_with_uci_to_transport_block_without_leaving_the_device(dec, tbs, Qm, layers, prb, mask, O_ack, re_ack, re_rvd,
                                                                        re_c1, re_c2, esn0):
    """A PUSCH that multiplexes HARQ-ACK / CSI with the transport block, received without the LLRs leaving HBM: equalised
    symbols -> demodulate + descramble (+ repetition placeholders) -> UL-SCH demultiplex -> rate dematch + decode + CRC ->
    transport block assembly. The transmitter is the oracle / numpy (the multiplexer is the oracle's demultiplexing walk
    inverted). Every soft bit of the demultiplexed fields equals the oracle chain and the transport block is recovered."""
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(tbs % 1000 + Qm)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    bpr = Qm * layers
    cfg = dict(mod_order=Qm, nof_layers=layers, nof_prb=prb, start_symbol=0, nof_symbols=14, dmrs_type=1, cdm=2, dmrs_mask=mask,
               rvd=re_rvd * bpr)
    G_ack, G_c1, G_c2 = re_ack * bpr, re_c1 * bpr, re_c2 * bpr
    ud = pkg.ulsch_desc(Qm, layers, prb, 0, 14, 1, 2, mask, cfg["rvd"], G_ack, G_c1, G_c2)
    n_in, n_sch = pkg.ulsch_sizes(ud)
    src = _source_map(cfg, G_ack, G_c1, G_c2, n_in)
    assert src[0].size == n_sch
    if re_rvd and O_ack:
        assert (src[0] < 0).any()  # the case really punctures shared-channel elements

    # transmitter
    tb, cws, meta = O.build_transport_block(rng, tbs, 1, Qm, layers, n_sch // Qm)
    descs, crc_bits = pkg.segment(tbs, 1, Qm, layers, n_sch // Qm)  # channel symbols over all layers
    C = descs.size
    tx = np.concatenate([O.port().rate_match(cws[r], int(meta[r, 2]), 0, Qm, 0, int(meta[r, 3])) for r in range(C)])
    assert tx.size == n_sch
    uci_bits = [rng.integers(0, 2, g).astype(np.uint8) for g in (G_ack, G_c1, G_c2)]
    mux = np.zeros(n_in, np.uint8)
    for field, bits in ((src[0], tx), (src[3], uci_bits[2]), (src[2], uci_bits[1]), (src[1], uci_bits[0])):
        keep = field >= 0
        mux[field[keep]] = bits[keep]  # HARQ-ACK last: it overwrites the elements it punctures
    c_init = (0x2345 << 15) + 99
    sym = _modulate(mux ^ O.port().gold_sequence(c_init, n_in), Qm)
    sigma2 = 10.0 ** (-esn0 / 10.0)
    sym = (sym + (rng.normal(0, 1, sym.size) + 1j * rng.normal(0, 1, sym.size)) * np.sqrt(sigma2 / 2)).astype(np.complex64)
    nv = np.full(sym.size, sigma2, np.float32)

    # receiver descriptors
    ro = so = oo = 0
    for r in range(C):
        d = descs[r]
        Z = int(d["lifting_size"])
        d["early_stop"], d["new_data"], d["in_offset"], d["soft_offset"], d["out_offset"] = 1, 1, ro, so, oo
        d["in_length"] = 66 * Z
        ro += int(d["rm_length"])
        so += 66 * 384
        oo += (22 * Z + 7) // 8
    assert ro == n_sch
    dec.softbuffer_reserve(so)
    job = np.zeros(1, pkg.DEMOD_DESC)
    job["mod_order"], job["descramble"], job["nof_symbols"], job["c_init"] = Qm, 1, sym.size, c_init
    pl = pkg.ulsch_placeholders(ud, O_ack, 4 if G_c1 else 0, 7 if G_c2 else 0)
    assert (pl.size > 0) == (O_ack == 1)
    ud["harq_ack_offset"], ud["csi_part1_offset"], ud["csi_part2_offset"] = 0, G_ack, G_ack + G_c1
    tbd = np.zeros(1, pkg.TB_DESC)
    tbd["nof_codeblocks"], tbd["tbs_bits"], tbd["nof_cb_crc_bits"] = C, tbs, int(crc_bits[0])
    sym_d = torch.from_numpy(sym.view(np.float32)).to(dev)
    nv_d = torch.from_numpy(nv).to(dev)
    llr_d = torch.zeros(n_in, dtype=torch.int8, device=dev)
    sch_d = torch.full((n_sch,), 55, dtype=torch.int8, device=dev)
    uci_d = torch.full((G_ack + G_c1 + G_c2 + 1,), 55, dtype=torch.int8, device=dev)
    out_d = torch.zeros(oo, dtype=torch.uint8, device=dev)
    res_d = torch.zeros(C, dtype=torch.int32, device=dev)
    tb_d = torch.zeros(tbs // 8 + 3, dtype=torch.uint8, device=dev)
    tbres_d = torch.zeros(2, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    # receiver: six enqueues on one stream, no host round trip in between
    dec.demodulate_batch_device(job, sym_d.data_ptr(), nv_d.data_ptr(), 0, llr_d.data_ptr())
    if pl.size:
        ph = np.zeros(1, pkg.PLACEHOLDER_DESC)
        ph["c_init"], ph["nof_llrs"], ph["nof_re"], ph["mod_order"], ph["nof_layers"] = c_init, n_in, pl.size, Qm, layers
        dec.descramble_placeholders_device(ph, pl, llr_d.data_ptr())
    dec.ulsch_demultiplex_batch_device(ud, llr_d.data_ptr(), sch_d.data_ptr(), uci_d.data_ptr())
    dec.dematch_decode_batch_device(descs, sch_d.data_ptr(), out_d.data_ptr(), res_d.data_ptr())
    dec.tb_assemble_device(tbd, descs, out_d.data_ptr(), res_d.data_ptr(), tb_d.data_ptr(), tbres_d.data_ptr())
    dec.synchronize()

    want_llr = O.port().demodulate_soft(Qm, sym, nv)
    want_llr = (O.port().descramble_placeholders(c_init, want_llr, Qm, layers, pl) if pl.size
                else O.port().descramble(c_init, want_llr))
    assert (llr_d.cpu().numpy() == want_llr).all()
    want = O.port().ulsch_demultiplex(cfg, G_ack, G_c1, G_c2, want_llr)
    assert (sch_d.cpu().numpy() == want[0]).all()
    uci = uci_d.cpu().numpy()
    assert (uci[:G_ack] == want[1]).all() and (uci[G_ack:G_ack + G_c1] == want[2]).all()
    assert (uci[G_ack + G_c1:G_ack + G_c1 + G_c2] == want[3]).all() and uci[-1] == 55
    # the control fields carry what was sent (hard decisions; the placeholder bits of a one-bit field are not information)
    if O_ack != 1:
        for off, bits in ((0, uci_bits[0]), (G_ack, uci_bits[1]), (G_ack + G_c1, uci_bits[2])):
            if bits.size:
                assert ((uci[off:off + bits.size] < 0) == (bits == 1)).mean() > 0.99
    tbres = tbres_d.cpu().numpy().view(pkg.TB_RESULT)
    assert int(tbres[0]["tb_crc_ok"]) == 1 and int(tbres[0]["nof_cb_failed"]) == 0
    assert (tb_d.cpu().numpy()[:tbs // 8] == tb).all()
