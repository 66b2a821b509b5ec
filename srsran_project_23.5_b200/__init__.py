"""B200 (sm_100a) PUSCH LDPC receive chain for srsRAN Project 23.5 - Python binding of the C ABI (include/ldpc_cuda.h).

The product is the shared library lib/libsrsran_ldpc_cuda.so (hand-written CUDA kernels + C ABI); this module is a thin
ctypes layer used by tests/ and bench.py. There is no CPU fallback anywhere: without the library, or without an sm_100
device, every compute call raises.

The directory name contains a dot, so it is loaded through ``__graft_entry__.load_package()`` (importlib) under the module
name ``srsran_project_23_5_b200``.
"""
from .ldpc import (CRC16, CRC24A, CRC24B, CRC_NONE, SCALE_AVX, SCALE_GENERIC, SCALE_NEON, FLAG_FORCE_SCALAR, FLAG_COPY_ONLY, CB_DESC,
                   CB_RESULT, TB_DESC, TB_RESULT, DEMOD_DESC, PLACEHOLDER_DESC, ULSCH_DESC, ulsch_desc, ulsch_sizes, ulsch_placeholders, LdpcCuda, LdpcCudaPool, LdpcCudaError, device_count, graph_shift, library, library_path, make_descs,
                   segment)

__all__ = ["CRC16", "CRC24A", "CRC24B", "CRC_NONE", "SCALE_AVX", "SCALE_GENERIC", "SCALE_NEON", "FLAG_FORCE_SCALAR", "FLAG_COPY_ONLY",
           "CB_DESC", "CB_RESULT", "TB_DESC", "TB_RESULT", "DEMOD_DESC", "PLACEHOLDER_DESC", "ULSCH_DESC", "ulsch_desc", "ulsch_sizes", "ulsch_placeholders", "LdpcCuda", "LdpcCudaPool", "LdpcCudaError", "device_count", "graph_shift", "library", "library_path",
           "make_descs", "segment"]
