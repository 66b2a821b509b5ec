"""GPU parity of the transmit-side kernels (LDPC encoder + rate matcher) against the oracle port, which
is itself pinned to the reference's encoder / rate matcher on every lifting size (tests/test_oracle_vs_reference.py), and
a loop-back through the receive chain: encode -> rate match -> noiseless LLRs -> fused dematch + decode."""
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


@pytest.mark.parametrize("bg", [1, 2])
def test_encoder_all_lifting_sizes(dec, bg):
    rng = np.random.default_rng(bg)
    for Z in O.LIFTING_SIZES:
        F = int(rng.integers(0, min(Z, 40))) if Z > 4 else 0
        msg = O.random_message(rng, bg, Z, F)
        want = O.port().encode(bg, Z, msg)
        got = dec.encode(bg, Z, msg)
        assert (got == want).all(), (bg, Z, np.nonzero(got != want)[0][:8])


def test_rate_matcher_cases(dec):
    from test_gpu_parity import RM_CASES
    rng = np.random.default_rng(5)
    for bg, Z, E, rv, Qm, Nref, F in RM_CASES:
        msg, cw = O.make_codeword(rng, bg, Z, F)
        want = O.port().rate_match(cw, E, rv, Qm, Nref, F)
        got = dec.rate_match(cw, E, bg, Z, rv, Qm, Nref, F)
        assert (got == want).all(), (bg, Z, E, rv, Qm, Nref, F)


def test_device_batch_loopback_through_the_receive_chain(dec):
    """Everything on the device: a batch of random messages -> encode -> rate match (rv0) -> +-20 LLRs -> fused
    dematch + decode with CRC-less fixed iterations -> the decoded bits equal the messages."""
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(9)
    shapes = [(1, 384, 16, 8, 9000), (1, 384, 0, 2, 12000), (2, 384, 0, 4, 8000), (1, 56, 12, 6, 1500), (2, 20, 0, 2, 400)]
    n = len(shapes)
    d_enc, d_rm, d_fus = np.zeros(n, pkg.CB_DESC), np.zeros(n, pkg.CB_DESC), np.zeros(n, pkg.CB_DESC)
    msgs, mo, co, ro, so, oo = [], 0, 0, 0, 0, 0
    for i, (bg, Z, F, Qm, E) in enumerate(shapes):
        K, N = O.BG_K[bg] * Z, O.BG_N_SHORT[bg] * Z
        msgs.append(O.random_message(rng, bg, Z, F))
        for d in (d_enc, d_rm, d_fus):
            d[i]["base_graph"], d[i]["lifting_size"], d[i]["nof_filler_bits"] = bg, Z, F
            d[i]["mod_order"], d[i]["rm_length"], d[i]["new_data"], d[i]["early_stop"] = Qm, E, 1, 1
        d_enc[i]["in_offset"], d_enc[i]["out_offset"] = mo, co
        d_rm[i]["in_offset"], d_rm[i]["out_offset"] = co, ro
        d_fus[i]["in_offset"], d_fus[i]["soft_offset"], d_fus[i]["out_offset"], d_fus[i]["in_length"] = ro, so, oo, N
        mo, co, ro, so, oo = mo + K, co + N, ro + E, so + N, oo + (K + 7) // 8
    msg_d = torch.from_numpy(np.concatenate(msgs)).to(dev)
    cw_d = torch.zeros(co, dtype=torch.uint8, device=dev)
    rm_d = torch.zeros(ro, dtype=torch.uint8, device=dev)
    stream = torch.cuda.ExternalStream(dec.stream, device=dev)
    torch.cuda.synchronize()
    dec.encode_batch_device(d_enc, msg_d.data_ptr(), cw_d.data_ptr())
    dec.rate_match_batch_device(d_rm, cw_d.data_ptr(), rm_d.data_ptr())
    with torch.cuda.stream(stream):
        llr_d = ((1 - 2 * rm_d.to(torch.int16)) * 20).to(torch.int8)
    out_d = torch.zeros(oo, dtype=torch.uint8, device=dev)
    res_d = torch.zeros(n, dtype=torch.int32, device=dev)
    dec.softbuffer_reserve(so)
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    stream.synchronize()
    dec.dematch_decode_batch_device(d_fus, llr_d.data_ptr(), out_d.data_ptr(), res_d.data_ptr())
    dec.synchronize()
    out = out_d.cpu().numpy()
    for i, (bg, Z, F, Qm, E) in enumerate(shapes):
        o = int(d_fus[i]["out_offset"])
        want = O.msg_to_packed(msgs[i])
        assert (out[o:o + want.size] == want).all(), i
