"""CPU-side checks of the product library: it loads, exports every symbol include/ldpc_cuda.h declares, its host-only
helpers agree with the oracle, and compute entry points fail loudly without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

import __graft_entry__ as ge
import oracles as O

pkg = ge.load_package()
ROOT = ge.ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "ldpc_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ldpc_cuda_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = pkg.library()
    names = declared_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), name
    assert lib.ldpc_cuda_abi_version() == 1


def test_struct_layout_matches_header():
    assert pkg.CB_DESC.itemsize == 48 and pkg.CB_RESULT.itemsize == 4
    assert pkg.CB_DESC.fields["in_offset"][1] == 24 and pkg.CB_DESC.fields["out_offset"][1] == 40
    assert pkg.PLACEHOLDER_DESC.itemsize == 24 and pkg.PLACEHOLDER_DESC.fields["nof_re"][1] == 20
    assert pkg.PLACEHOLDER_DESC.fields["mod_order"][1] == 22 and pkg.PLACEHOLDER_DESC.fields["nof_layers"][1] == 23


def test_no_product_source_touches_the_oracle():
    """The product path must never route through oracle/ (or any CPU fallback)."""
    for base, _, files in os.walk(os.path.join(ROOT, "srsran_project_23.5_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                text = open(os.path.join(base, f), errors="ignore").read()
                assert "oracle" not in text.replace("oracle/ (", ""), os.path.join(base, f)


@pytest.mark.skipif(pkg.device_count() > 0, reason="only meaningful on a machine without an sm_100 GPU")
def test_compute_fails_loudly_without_gpu():
    with pytest.raises(pkg.LdpcCudaError) as e:
        pkg.LdpcCuda(device=0)
    assert e.value.status == -2


def test_graph_shift_helper_all_graphs():
    p = O.port()
    for bg in (1, 2):
        for Z in O.LIFTING_SIZES:
            M, N = O.BG_M[bg], O.BG_K[bg] + O.BG_M[bg]
            for m in (0, 1, 3, 4, 17, M - 1):
                for n in range(N):
                    assert pkg.graph_shift(bg, Z, m, n) == p.lib.oracle_graph_shift(bg, Z, m, n)
    assert pkg.graph_shift(1, 17, 0, 0) == -1 and pkg.graph_shift(3, 16, 0, 0) == -1


def test_segment_helper_matches_oracle():
    p = O.port()
    rng = np.random.default_rng(0)
    cases = [(1, 8, 1, 100, 1), (1, 3824, 2, 4000, 1), (1, 3832, 2, 4000, 1), (1, 8424, 4, 6000, 2),
             (1, 319784, 8, 42588, 1), (1, 1277992, 8, 170352, 4), (2, 24, 2, 144, 1), (2, 3840, 4, 2400, 2),
             (2, 30000, 6, 12000, 4)]
    for _ in range(40):
        bg = int(rng.integers(1, 3))
        nl = int(rng.integers(1, 5))
        cases.append((bg, int(rng.integers(3, 60000 if bg == 2 else 400000)) * 8, int(rng.choice([1, 2, 4, 6, 8])),
                      nl * int(rng.integers(50, 30000)), nl))
    for bg, tbs, Qm, nsym, nl in cases:
        want = p.segment_rx(tbs, bg, Qm, nl, nsym)
        descs, crc_bits = pkg.segment(tbs, bg, Qm, nl, nsym, rv=2, Nref=0)
        assert descs.size == want.shape[0]
        assert (descs["lifting_size"] == want[:, 0]).all() and (descs["rm_length"] == want[:, 2]).all()
        assert (descs["nof_filler_bits"] == want[:, 3]).all() and (descs["in_offset"] == want[:, 4]).all()
        assert (crc_bits == want[:, 5]).all() and (descs["rv"] == 2).all()
        C = descs.size
        crc = pkg.CRC24B if C > 1 else (pkg.CRC24A if tbs > 3824 else pkg.CRC16)  # select_crc
        assert (descs["crc_poly"] == crc).all()


def test_invalid_arguments_rejected_by_segment():
    lib = pkg.library()
    assert lib.ldpc_cuda_segment(1000, 3, 0, 2, 0, 1, 100, None, 0, None) == -1
    assert lib.ldpc_cuda_segment(1000, 1, 0, 3, 0, 1, 100, None, 0, None) == -1
    assert lib.ldpc_cuda_segment(1000, 1, 0, 2, 0, 2, 101, None, 0, None) == -1
