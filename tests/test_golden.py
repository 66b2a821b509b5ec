"""Known-answer tests against tests/golden/ldpc_golden.npz, vectors produced by the reference's own code
(tests/golden/make_golden.py; the reference's MATLAB tarballs are not in its tree). The CPU half checks the oracle port,
the gpu half checks the CUDA path through the C ABI. /root/reference is not needed at test time."""
import os

import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()
G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ldpc_golden.npz"))
N_DEC = sum(1 for k in G.files if k.endswith("_llr"))
N_RM = sum(1 for k in G.files if k.endswith("_prior"))


def test_fixture_is_complete():
    assert N_DEC >= 12 and N_RM >= 10 and "crc_vals" in G.files


@pytest.mark.parametrize("i", range(N_DEC))
def test_oracle_decoder_matches_golden(i):
    bg, Z, F, crc, iters, it = [int(x) for x in G["dec%d_meta" % i]]
    bits, got_it = O.port().decode(bg, Z, G["dec%d_llr" % i], fillers=F, crc=crc, max_iters=iters)
    assert got_it == it and (bits == G["dec%d_bits" % i]).all()
    if "dec%d_avx2_bits" % i in G.files:
        bits, got_it = O.port().decode(bg, Z, G["dec%d_llr" % i], fillers=F, crc=crc, max_iters=iters,
                                       scale_mode=O.SCALE_AVX)
        assert got_it == int(G["dec%d_avx2_it" % i][0]) and (bits == G["dec%d_avx2_bits" % i]).all()


@pytest.mark.parametrize("i", range(N_RM))
def test_oracle_dematcher_matches_golden(i):
    bg, Z, E, rv, Qm, Nref, F, retx = [int(x) for x in G["rm%d_meta" % i]]
    soft = G["rm%d_prior" % i].copy()
    for t in range(1 + retx):
        O.port().rate_dematch(soft, G["rm%d_in%d" % (i, t)], t == 0, rv, Qm, Nref, F)
        assert (soft == G["rm%d_out%d" % (i, t)]).all(), (i, t)


def test_oracle_crc_matches_golden():
    for p, poly in enumerate((O.CRC16, O.CRC24A, O.CRC24B)):
        for k, n in enumerate(G["crc_lens"]):
            assert O.port().crc(poly, G["crc_data"], int(n)) == int(G["crc_vals"][p, k])


@pytest.fixture(scope="module", params=[0, pkg.FLAG_FORCE_SCALAR], ids=["auto", "scalar"])
def dec(request):
    d = pkg.LdpcCuda(device=0, max_codeblocks=64, flags=request.param)
    yield d
    d.close()


@pytest.mark.gpu
@pytest.mark.parametrize("i", range(N_DEC))
def test_cuda_decoder_matches_golden(dec, i):
    bg, Z, F, crc, iters, it = [int(x) for x in G["dec%d_meta" % i]]
    dec.set_algorithm(iters, 0.8, O.SCALE_GENERIC)
    bits, got_it = dec.decode(bg, Z, G["dec%d_llr" % i], fillers=F, crc=crc)
    assert got_it == it and (bits == G["dec%d_bits" % i]).all()
    if "dec%d_avx2_bits" % i in G.files:
        dec.set_algorithm(iters, 0.8, O.SCALE_AVX)
        bits, got_it = dec.decode(bg, Z, G["dec%d_llr" % i], fillers=F, crc=crc)
        assert got_it == int(G["dec%d_avx2_it" % i][0]) and (bits == G["dec%d_avx2_bits" % i]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("i", range(N_RM))
def test_cuda_dematcher_matches_golden(dec, i):
    bg, Z, E, rv, Qm, Nref, F, retx = [int(x) for x in G["rm%d_meta" % i]]
    soft = G["rm%d_prior" % i].copy()
    for t in range(1 + retx):
        dec.rate_dematch(soft, G["rm%d_in%d" % (i, t)], t == 0, bg, Z, rv, Qm, Nref, F)
        assert (soft == G["rm%d_out%d" % (i, t)]).all(), (i, t)


@pytest.mark.gpu
def test_cuda_batch_of_all_golden_codeblocks(dec):
    """All golden decoder cases with the default 6 iterations in ONE heterogeneous batched call."""
    dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
    idx = [i for i in range(N_DEC) if int(G["dec%d_meta" % i][4]) == 6]
    descs, llrs, in_off, out_off = [], [], 0, 0
    for i in idx:
        bg, Z, F, crc, iters, it = [int(x) for x in G["dec%d_meta" % i]]
        x = G["dec%d_llr" % i]
        d = pkg.make_descs(1, bg, Z, in_length=x.size, nof_filler_bits=F, crc_poly=crc)
        d["in_offset"], d["out_offset"] = in_off, out_off
        in_off += x.size
        out_off += (O.BG_K[bg] * Z + 7) // 8
        descs.append(d)
        llrs.append(x)
    descs = np.concatenate(descs)
    out, res = dec.decode_batch(descs, np.concatenate(llrs))
    for k, i in enumerate(idx):
        it = int(G["dec%d_meta" % i][5])
        bits = G["dec%d_bits" % i]
        o = int(descs["out_offset"][k])
        assert (out[o:o + bits.size] == bits).all(), i
        assert (int(res["iterations"][k]) if res["crc_ok"][k] else 0) == it, i
