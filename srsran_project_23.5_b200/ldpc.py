"""ctypes binding of include/ldpc_cuda.h. Names and argument meaning follow the reference interfaces:
ldpc_decoder::decode (ldpc_decoder.h:44-74), ldpc_rate_dematcher::rate_dematch (ldpc_rate_dematcher.h:52-55),
codeblock_metadata (codeblock_metadata.h:38-73). Buffers are numpy arrays (host entries) or raw device pointers
(device entries, e.g. ``torch.Tensor.data_ptr()``)."""
import ctypes
import os

import numpy as np

CRC_NONE, CRC16, CRC24A, CRC24B = 0, 1, 2, 3
SCALE_GENERIC, SCALE_AVX, SCALE_NEON = 0, 1, 2
FLAG_FORCE_SCALAR = 1
FLAG_COPY_ONLY = 2

NOALLOC_MAX_TBS, NOALLOC_MAX_CBS, NOALLOC_MAX_CHUNKS = 64, 4096, 8192  # ldpc_cuda_tb_assemble_device_noalloc

BG_K = {1: 22, 2: 10}
BG_N_SHORT = {1: 66, 2: 50}

# numpy mirror of ldpc_cuda_cb_desc (48 bytes) and ldpc_cuda_cb_result (4 bytes)
CB_DESC = np.dtype([("base_graph", "u1"), ("crc_poly", "u1"), ("rv", "u1"), ("mod_order", "u1"),
                    ("lifting_size", "<u2"), ("nof_filler_bits", "<u2"), ("Nref", "<u4"), ("rm_length", "<u4"),
                    ("in_length", "<u4"), ("new_data", "u1"), ("skip_decode", "u1"), ("early_stop", "u1"),
                    ("reserved", "u1"), ("in_offset", "<u8"), ("soft_offset", "<u8"), ("out_offset", "<u8")],
                   align=True)
CB_RESULT = np.dtype([("iterations", "<u2"), ("crc_ok", "u1"), ("flags", "u1")], align=True)
assert CB_DESC.itemsize == 48 and CB_RESULT.itemsize == 4
# ldpc_cuda_tb_desc (24 bytes) and ldpc_cuda_tb_result (8 bytes)
TB_DESC = np.dtype([("first_codeblock", "<u4"), ("nof_codeblocks", "<u4"), ("tbs_bits", "<u4"),
                    ("nof_cb_crc_bits", "<u4"), ("tb_offset", "<u8")], align=True)
TB_RESULT = np.dtype([("crc_remainder", "<u4"), ("nof_cb_failed", "<u2"), ("tb_crc_ok", "u1"), ("reserved", "u1")],
                     align=True)
assert TB_DESC.itemsize == 24 and TB_RESULT.itemsize == 8
# ldpc_cuda_demod_desc (32 bytes)
DEMOD_DESC = np.dtype([("mod_order", "u1"), ("pi2_bpsk", "u1"), ("descramble", "u1"), ("reserved", "u1"),
                       ("nof_symbols", "<u4"), ("c_init", "<u4"), ("reserved2", "<u4"), ("symbol_offset", "<u8"),
                       ("llr_offset", "<u8")], align=True)
assert DEMOD_DESC.itemsize == 32
# ldpc_cuda_ulsch_desc (72 bytes)
ULSCH_DESC = np.dtype([("mod_order", "u1"), ("nof_layers", "u1"), ("start_symbol_index", "u1"), ("nof_symbols", "u1"),
                       ("dmrs_type", "u1"), ("nof_cdm_groups_without_data", "u1"), ("dmrs_symbol_mask", "<u2"),
                       ("nof_prb", "<u2"), ("reserved", "<u2"), ("nof_harq_ack_rvd", "<u4"),
                       ("nof_enc_harq_ack_bits", "<u4"), ("nof_enc_csi_part1_bits", "<u4"),
                       ("nof_enc_csi_part2_bits", "<u4"), ("in_offset", "<u8"), ("sch_offset", "<u8"),
                       ("harq_ack_offset", "<u8"), ("csi_part1_offset", "<u8"), ("csi_part2_offset", "<u8")], align=True)
assert ULSCH_DESC.itemsize == 72
# ldpc_cuda_placeholder_desc (24 bytes)
PLACEHOLDER_DESC = np.dtype([("llr_offset", "<u8"), ("c_init", "<u4"), ("nof_llrs", "<u4"), ("first_re", "<u4"),
                             ("nof_re", "<u2"), ("mod_order", "u1"), ("nof_layers", "u1")], align=True)
assert PLACEHOLDER_DESC.itemsize == 24


class LdpcCudaError(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        super().__init__("%s: %s (status %d)" % (where, library().ldpc_cuda_status_string(status).decode(), status))


class _Algorithm(ctypes.Structure):
    _fields_ = [("max_iterations", ctypes.c_uint32), ("scaling_factor", ctypes.c_float),
                ("scale_mode", ctypes.c_uint32), ("reserved", ctypes.c_uint32)]


_lib = None


def library_path():
    override = os.environ.get("SRSRAN_LDPC_CUDA_LIB")  # development only: A/B runs of two builds of the same library
    if override:
        return override
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libsrsran_ldpc_cuda.so")


def library():
    """Loads the CUDA extension. Fails loudly when it has not been built - there is no fallback path."""
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            raise ImportError("CUDA extension %s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)" % path)
        L = ctypes.CDLL(path)
        vp, u32, u64, i32 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_int
        L.ldpc_cuda_status_string.restype = ctypes.c_char_p
        L.ldpc_cuda_status_string.argtypes = [i32]
        L.ldpc_cuda_create.argtypes = [ctypes.POINTER(vp), i32, u32, u32]
        L.ldpc_cuda_destroy.argtypes = [vp]
        L.ldpc_cuda_destroy.restype = None
        L.ldpc_cuda_set_algorithm.argtypes = [vp, ctypes.POINTER(_Algorithm)]
        L.ldpc_cuda_stream.argtypes = [vp]
        L.ldpc_cuda_stream.restype = vp
        L.ldpc_cuda_synchronize.argtypes = [vp]
        L.ldpc_cuda_launch_count.argtypes = [vp]
        L.ldpc_cuda_launch_count.restype = u64
        L.ldpc_cuda_general_kernel_launch_count.argtypes = [vp]
        L.ldpc_cuda_general_kernel_launch_count.restype = u64
        L.ldpc_cuda_decode.argtypes = [vp, vp, vp, vp, ctypes.POINTER(u32)]
        L.ldpc_cuda_rate_dematch.argtypes = [vp, vp, vp, vp]
        for name in ("ldpc_cuda_decode_batch", "ldpc_cuda_decode_batch_device", "ldpc_cuda_dematch_decode_batch",
                     "ldpc_cuda_dematch_decode_batch_device"):
            getattr(L, name).argtypes = [vp, u32, vp, vp, vp, vp]
        L.ldpc_cuda_softbuffer_reserve.argtypes = [vp, u64]
        L.ldpc_cuda_softbuffer_zero.argtypes = [vp, u64, u64]
        L.ldpc_cuda_softbuffer_read.argtypes = [vp, u64, vp, u64]
        L.ldpc_cuda_softbuffer_write.argtypes = [vp, u64, vp, u64]
        L.ldpc_cuda_softbuffer_ptr.argtypes = [vp]
        L.ldpc_cuda_softbuffer_ptr.restype = vp
        L.ldpc_cuda_arena_create.argtypes = [i32, u64, ctypes.POINTER(vp)]
        L.ldpc_cuda_arena_destroy.argtypes = [i32, vp]
        L.ldpc_cuda_softbuffer_attach.argtypes = [vp, vp, u64]
        for name in ("ldpc_cuda_encode_batch_device", "ldpc_cuda_rate_match_batch_device"):
            getattr(L, name).argtypes = [vp, u32, vp, vp, vp]
        for name in ("ldpc_cuda_encode", "ldpc_cuda_rate_match"):
            getattr(L, name).argtypes = [vp, vp, vp, vp]
        L.ldpc_cuda_tb_assemble_device.argtypes = [vp, u32, vp, u32, vp, vp, vp, vp, vp]
        L.ldpc_cuda_tb_assemble_device_noalloc.argtypes = [vp, u32, vp, u32, vp, vp, vp, vp, vp]
        L.ldpc_cuda_dematch_decode_tb_batch.argtypes = [vp, u32, vp, vp, vp, vp, u32, vp, vp, vp]
        L.ldpc_cuda_demodulate_batch_device.argtypes = [vp, u32, vp, vp, vp, vp, vp]
        L.ldpc_cuda_descramble_placeholders_device.argtypes = [vp, u32, vp, vp, u32, vp]
        L.ldpc_cuda_descramble_uci.argtypes = [vp, u32, u32, ctypes.c_int, u32, u32, vp, vp, vp]
        L.ldpc_cuda_demodulate_soft.argtypes = [vp, i32, i32, u32, vp, vp, vp]
        L.ldpc_cuda_descramble.argtypes = [vp, u32, u32, vp, vp]
        L.ldpc_cuda_segment.argtypes = [u32, i32, i32, i32, u32, u32, u32, vp, u32, vp]
        L.ldpc_cuda_graph_shift.argtypes = [i32, i32, i32, i32]
        L.ldpc_cuda_ulsch_sizes.argtypes = [vp, ctypes.POINTER(u32), ctypes.POINTER(u32)]
        L.ldpc_cuda_ulsch_demultiplex_batch_device.argtypes = [vp, u32, vp, vp, vp, vp]
        L.ldpc_cuda_ulsch_demultiplex.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.ldpc_cuda_ulsch_placeholders.argtypes = [vp, u32, u32, u32, vp, u32]
        L.ldpc_cuda_pool_create.argtypes = [ctypes.POINTER(vp), ctypes.POINTER(i32), u32, u32, u32]
        L.ldpc_cuda_pool_destroy.argtypes = [vp]
        L.ldpc_cuda_pool_destroy.restype = None
        L.ldpc_cuda_pool_nof_devices.argtypes = [vp]
        L.ldpc_cuda_pool_nof_devices.restype = u32
        L.ldpc_cuda_pool_handle.argtypes = [vp, u32]
        L.ldpc_cuda_pool_handle.restype = vp
        L.ldpc_cuda_pool_set_algorithm.argtypes = [vp, ctypes.POINTER(_Algorithm)]
        L.ldpc_cuda_pool_softbuffer_reserve.argtypes = [vp, u64]
        L.ldpc_cuda_pool_launch_count.argtypes = [vp]
        L.ldpc_cuda_pool_launch_count.restype = u64
        L.ldpc_cuda_pool_codeblocks_per_device.argtypes = [vp, vp, u32]
        L.ldpc_cuda_pool_device_of.argtypes = [vp, u64]
        L.ldpc_cuda_pool_decode_batch.argtypes = [vp, u32, vp, vp, vp, vp]
        L.ldpc_cuda_pool_decode_batch_device.argtypes = [vp, u32, vp, vp, vp, vp]
        L.ldpc_cuda_pool_synchronize.argtypes = [vp]
        L.ldpc_cuda_pool_dematch_decode_tb_batch.argtypes = [vp, u32, vp, vp, vp, vp, u32, vp, vp, vp, vp]
        _lib = L
    return _lib


def device_count():
    return library().ldpc_cuda_device_count()


def graph_shift(bg, Z, m, n):
    return library().ldpc_cuda_graph_shift(bg, Z, m, n)


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def segment(tbs, bg, mod_order, nof_layers, nof_ch_symbols, rv=0, Nref=0):
    """ldpc_segmenter_rx::segment geometry. Returns (descs[C], nof_crc_bits[C])."""
    L = library()
    C = L.ldpc_cuda_segment(tbs, bg, rv, mod_order, Nref, nof_layers, nof_ch_symbols, None, 0, None)
    if C <= 0:
        raise LdpcCudaError(C, "ldpc_cuda_segment")
    descs = np.zeros(C, CB_DESC)
    crc_bits = np.zeros(C, np.uint32)
    L.ldpc_cuda_segment(tbs, bg, rv, mod_order, Nref, nof_layers, nof_ch_symbols, _ptr(descs), C, _ptr(crc_bits))
    return descs, crc_bits


def ulsch_desc(mod_order, nof_layers, nof_prb, start_symbol, nof_symbols, dmrs_type, cdm, dmrs_mask, rvd, G_ack=0, G_csi1=0,
               G_csi2=0):
    """One ldpc_cuda_ulsch_desc (ulsch_demultiplex::configuration + the sizes of the three UCI fields)."""
    d = np.zeros(1, ULSCH_DESC)
    d["mod_order"], d["nof_layers"], d["nof_prb"], d["start_symbol_index"] = mod_order, nof_layers, nof_prb, start_symbol
    d["nof_symbols"], d["dmrs_type"], d["nof_cdm_groups_without_data"], d["dmrs_symbol_mask"] = nof_symbols, dmrs_type, cdm, dmrs_mask
    d["nof_harq_ack_rvd"], d["nof_enc_harq_ack_bits"] = rvd, G_ack
    d["nof_enc_csi_part1_bits"], d["nof_enc_csi_part2_bits"] = G_csi1, G_csi2
    return d


def ulsch_sizes(desc):
    """(input soft bits, SCH soft bits) of a descriptor, or None for a configuration the reference asserts on."""
    a, b = ctypes.c_uint32(0), ctypes.c_uint32(0)
    rc = library().ldpc_cuda_ulsch_sizes(_ptr(desc), ctypes.byref(a), ctypes.byref(b))
    return None if rc != 0 else (a.value, b.value)


def ulsch_placeholders(desc, O_ack, O_csi1, O_csi2):
    """ulsch_demultiplex::get_placeholders (host only): RE indexes, or None for an inconsistent configuration."""
    lst = np.zeros(1 << 16, np.uint16)
    n = library().ldpc_cuda_ulsch_placeholders(_ptr(desc), O_ack, O_csi1, O_csi2, _ptr(lst), lst.size)
    return None if n < 0 else lst[:n].copy()


def make_descs(n, bg, Z, **fields):
    """n descriptors of one (base graph, lifting size) with dense default offsets for decode-only batches."""
    d = np.zeros(n, CB_DESC)
    d["base_graph"], d["lifting_size"] = bg, Z
    d["mod_order"] = 1
    d["early_stop"] = 1
    d["new_data"] = 1
    in_len = fields.get("in_length", BG_N_SHORT[bg] * Z)
    d["in_length"] = in_len
    out_bytes = (BG_K[bg] * Z + 7) // 8
    d["in_offset"] = np.arange(n, dtype=np.uint64) * np.uint64(fields.pop("in_stride", in_len))
    d["out_offset"] = np.arange(n, dtype=np.uint64) * np.uint64(fields.pop("out_stride", out_bytes))
    for k, v in fields.items():
        d[k] = v
    return d


class LdpcCuda:
    """One handle = one GPU stream + scratch, like one reference decoder/dematcher object (not thread-safe)."""

    def __init__(self, device=0, max_codeblocks=1024, flags=0, max_iterations=6, scaling_factor=0.8,
                 scale_mode=SCALE_GENERIC):
        self._L = library()
        h = ctypes.c_void_p()
        rc = self._L.ldpc_cuda_create(ctypes.byref(h), device, max_codeblocks, flags)
        if rc != 0:
            raise LdpcCudaError(rc, "ldpc_cuda_create")
        self._h = h
        self.device = device
        self.flags = flags
        self.set_algorithm(max_iterations, scaling_factor, scale_mode)

    def close(self):
        if getattr(self, "_h", None):
            self._L.ldpc_cuda_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, where):
        if rc != 0:
            raise LdpcCudaError(rc, where)

    def set_algorithm(self, max_iterations=6, scaling_factor=0.8, scale_mode=SCALE_GENERIC):
        alg = _Algorithm(max_iterations, scaling_factor, scale_mode, 0)
        self._check(self._L.ldpc_cuda_set_algorithm(self._h, ctypes.byref(alg)), "ldpc_cuda_set_algorithm")
        self.max_iterations = max_iterations

    @property
    def stream(self):
        return self._L.ldpc_cuda_stream(self._h)

    def synchronize(self):
        self._check(self._L.ldpc_cuda_synchronize(self._h), "ldpc_cuda_synchronize")

    @property
    def launch_count(self):
        return self._L.ldpc_cuda_launch_count(self._h)

    @property
    def general_kernel_launch_count(self):
        return self._L.ldpc_cuda_general_kernel_launch_count(self._h)

    # -- drop-in, one codeblock ----------------------------------------------------------------------------------------
    def decode(self, bg, Z, llrs, fillers=0, crc=CRC_NONE, out=None):
        """ldpc_decoder::decode. Returns (packed bits, iterations) with iterations == 0 <-> nullopt."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        d = make_descs(1, bg, Z, in_length=llrs.size, nof_filler_bits=fillers, crc_poly=crc)
        if out is None:
            out = np.zeros((BG_K[bg] * Z + 7) // 8, np.uint8)
        it = ctypes.c_uint32(0)
        self._check(self._L.ldpc_cuda_decode(self._h, _ptr(d), _ptr(llrs), _ptr(out), ctypes.byref(it)),
                    "ldpc_cuda_decode")
        return out, it.value

    def rate_dematch(self, soft, llrs, new_data, bg, Z, rv=0, mod_order=1, Nref=0, fillers=0):
        """ldpc_rate_dematcher::rate_dematch; soft (int8[N_short*Z]) is updated in place."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        assert soft.dtype == np.int8 and soft.flags.c_contiguous and soft.size == BG_N_SHORT[bg] * Z
        d = make_descs(1, bg, Z, rm_length=llrs.size, rv=rv, mod_order=mod_order, Nref=Nref, nof_filler_bits=fillers,
                       new_data=int(bool(new_data)))
        self._check(self._L.ldpc_cuda_rate_dematch(self._h, _ptr(d), _ptr(soft), _ptr(llrs)), "ldpc_cuda_rate_dematch")
        return soft

    # -- batched -------------------------------------------------------------------------------------------------------
    def decode_batch(self, descs, llrs, out=None, results=None):
        descs = np.ascontiguousarray(descs, CB_DESC)
        llrs = np.ascontiguousarray(llrs, np.int8)
        n = descs.size
        if out is None:
            span = int((descs["out_offset"] + (descs["lifting_size"].astype(np.uint64) *
                                               np.where(descs["base_graph"] == 2, 10, 22).astype(np.uint64) + 7) // 8
                        ).max()) if n else 0
            out = np.zeros(span, np.uint8)
        if results is None:
            results = np.zeros(n, CB_RESULT)
        self._check(self._L.ldpc_cuda_decode_batch(self._h, n, _ptr(descs), _ptr(llrs), _ptr(out), _ptr(results)),
                    "ldpc_cuda_decode_batch")
        return out, results

    def decode_batch_device(self, descs, llrs_ptr, out_ptr, results_ptr):
        descs = np.ascontiguousarray(descs, CB_DESC)
        self._check(self._L.ldpc_cuda_decode_batch_device(self._h, descs.size, _ptr(descs), llrs_ptr, out_ptr,
                                                          results_ptr), "ldpc_cuda_decode_batch_device")

    def dematch_decode_batch(self, descs, rm_llrs, out, results=None):
        descs = np.ascontiguousarray(descs, CB_DESC)
        rm_llrs = np.ascontiguousarray(rm_llrs, np.int8)
        if results is None:
            results = np.zeros(descs.size, CB_RESULT)
        self._check(self._L.ldpc_cuda_dematch_decode_batch(self._h, descs.size, _ptr(descs), _ptr(rm_llrs), _ptr(out),
                                                           _ptr(results)), "ldpc_cuda_dematch_decode_batch")
        return out, results

    def dematch_decode_batch_device(self, descs, rm_ptr, out_ptr, results_ptr):
        descs = np.ascontiguousarray(descs, CB_DESC)
        self._check(self._L.ldpc_cuda_dematch_decode_batch_device(self._h, descs.size, _ptr(descs), rm_ptr, out_ptr,
                                                                  results_ptr), "ldpc_cuda_dematch_decode_batch_device")

    # -- transport-block assembly + TB CRC24A on the device (pusch_decoder_impl.cpp:156-222) ----------------------------
    def tb_assemble_device(self, tbs, descs, out_ptr, results_ptr, tb_out_ptr, tb_results_ptr):
        tbs = np.ascontiguousarray(tbs, TB_DESC)
        descs = np.ascontiguousarray(descs, CB_DESC)
        self._check(self._L.ldpc_cuda_tb_assemble_device(self._h, tbs.size, _ptr(tbs), descs.size, _ptr(descs), out_ptr,
                                                         results_ptr, tb_out_ptr, tb_results_ptr),
                    "ldpc_cuda_tb_assemble_device")

    def tb_assemble_device_noalloc(self, tbs, descs, out_ptr, results_ptr, tb_out_ptr, tb_results_ptr):
        """tb_assemble_device on a fixed device array: no allocation in the call (limits: NOALLOC_MAX_*)."""
        tbs = np.ascontiguousarray(tbs, TB_DESC)
        descs = np.ascontiguousarray(descs, CB_DESC)
        self._check(self._L.ldpc_cuda_tb_assemble_device_noalloc(self._h, tbs.size, _ptr(tbs), descs.size, _ptr(descs),
                                                                 out_ptr, results_ptr, tb_out_ptr, tb_results_ptr),
                    "ldpc_cuda_tb_assemble_device_noalloc")

    def dematch_decode_tb_batch(self, descs, rm_llrs, out, tbs, tb_out, results=None, tb_results=None):
        """out=None: the codeblock messages stay on the device (no skip_decode codeblocks allowed)."""
        descs = np.ascontiguousarray(descs, CB_DESC)
        tbs = np.ascontiguousarray(tbs, TB_DESC)
        rm_llrs = np.ascontiguousarray(rm_llrs, np.int8)
        if results is None:
            results = np.zeros(descs.size, CB_RESULT)
        if tb_results is None:
            tb_results = np.zeros(tbs.size, TB_RESULT)
        self._check(self._L.ldpc_cuda_dematch_decode_tb_batch(self._h, descs.size, _ptr(descs), _ptr(rm_llrs),
                                                              None if out is None else _ptr(out),
                                                              _ptr(results), tbs.size, _ptr(tbs), _ptr(tb_out),
                                                              _ptr(tb_results)), "ldpc_cuda_dematch_decode_tb_batch")
        return out, results, tb_out, tb_results

    # -- soft demodulator + descrambler (demodulation_mapper::demodulate_soft, pusch_demodulator_impl::descramble) ------
    def demodulate_soft(self, mod_order, symbols, noise_vars, pi2=False):
        """symbols: complex64 [n], noise_vars: float32 [n] -> int8 [n * mod_order]."""
        symbols = np.ascontiguousarray(symbols, np.complex64)
        noise_vars = np.ascontiguousarray(noise_vars, np.float32)
        out = np.zeros(symbols.size * mod_order, np.int8)
        self._check(self._L.ldpc_cuda_demodulate_soft(self._h, mod_order, int(pi2), symbols.size, _ptr(symbols),
                                                      _ptr(noise_vars), _ptr(out)), "ldpc_cuda_demodulate_soft")
        return out

    def descramble(self, c_init, llrs):
        llrs = np.ascontiguousarray(llrs, np.int8)
        out = np.zeros_like(llrs)
        self._check(self._L.ldpc_cuda_descramble(self._h, c_init, llrs.size, _ptr(llrs), _ptr(out)), "ldpc_cuda_descramble")
        return out

    def descramble_uci(self, c_init, llrs, mod_order, nof_layers, re_indexes):
        """pusch_demodulator_impl::descramble with the repetition placeholders of the listed REs, host buffers."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        re_indexes = np.ascontiguousarray(re_indexes, np.uint16)
        out = np.zeros_like(llrs)
        self._check(self._L.ldpc_cuda_descramble_uci(self._h, c_init, llrs.size, mod_order, nof_layers, re_indexes.size,
                                                     _ptr(re_indexes), _ptr(llrs), _ptr(out)), "ldpc_cuda_descramble_uci")
        return out

    def ulsch_demultiplex(self, desc, llrs):
        """ulsch_demultiplex::demultiplex, host buffers. Returns (sch_data, harq_ack, csi_part1, csi_part2)."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        sizes = ulsch_sizes(desc)
        if sizes is None or sizes[0] != llrs.size:
            raise LdpcCudaError(-1, "ldpc_cuda_ulsch_sizes")
        sch = np.zeros(sizes[1], np.int8)
        ack = np.zeros(int(desc["nof_enc_harq_ack_bits"][0]), np.int8)
        c1 = np.zeros(int(desc["nof_enc_csi_part1_bits"][0]), np.int8)
        c2 = np.zeros(int(desc["nof_enc_csi_part2_bits"][0]), np.int8)
        self._check(self._L.ldpc_cuda_ulsch_demultiplex(self._h, _ptr(desc), _ptr(llrs), _ptr(sch),
                                                        _ptr(ack) if ack.size else None, _ptr(c1) if c1.size else None,
                                                        _ptr(c2) if c2.size else None), "ldpc_cuda_ulsch_demultiplex")
        return sch, ack, c1, c2

    def ulsch_demultiplex_batch_device(self, descs, in_ptr, sch_ptr, uci_ptr):
        descs = np.ascontiguousarray(descs, ULSCH_DESC)
        self._check(self._L.ldpc_cuda_ulsch_demultiplex_batch_device(self._h, descs.size, _ptr(descs), in_ptr, sch_ptr,
                                                                     uci_ptr), "ldpc_cuda_ulsch_demultiplex_batch_device")

    def demodulate_batch_device(self, jobs, symbols_ptr, noise_ptr, in_llrs_ptr, llrs_ptr):
        jobs = np.ascontiguousarray(jobs, DEMOD_DESC)
        self._check(self._L.ldpc_cuda_demodulate_batch_device(self._h, jobs.size, _ptr(jobs), symbols_ptr, noise_ptr,
                                                              in_llrs_ptr, llrs_ptr), "ldpc_cuda_demodulate_batch_device")

    def descramble_placeholders_device(self, jobs, re_indexes, llrs_ptr):
        """UCI repetition placeholders (pusch_demodulator_impl.cpp:100-152) applied to codewords in HBM that were
        descrambled as plain data; asynchronous on the handle's stream."""
        jobs = np.ascontiguousarray(jobs, PLACEHOLDER_DESC)
        re_indexes = np.ascontiguousarray(re_indexes, np.uint16)
        self._check(self._L.ldpc_cuda_descramble_placeholders_device(self._h, jobs.size, _ptr(jobs), _ptr(re_indexes),
                                                                     re_indexes.size, llrs_ptr),
                    "ldpc_cuda_descramble_placeholders_device")

    # -- transmit side (encoder + rate matcher) ------------------------------------------------------------------------
    def encode(self, bg, Z, msg_bits):
        """ldpc_encoder::encode: bg_K*Z unpacked bits (fillers = 254) -> N_short*Z codeblock bits."""
        msg_bits = np.ascontiguousarray(msg_bits, np.uint8)
        assert msg_bits.size == BG_K[bg] * Z
        d = make_descs(1, bg, Z)
        out = np.zeros(BG_N_SHORT[bg] * Z, np.uint8)
        self._check(self._L.ldpc_cuda_encode(self._h, _ptr(d), _ptr(msg_bits), _ptr(out)), "ldpc_cuda_encode")
        return out

    def rate_match(self, codeblock, E, bg, Z, rv=0, mod_order=1, Nref=0, fillers=0):
        """ldpc_rate_matcher::rate_match: N_short*Z codeblock bits -> E rate-matched, interleaved bits."""
        codeblock = np.ascontiguousarray(codeblock, np.uint8)
        assert codeblock.size == BG_N_SHORT[bg] * Z
        d = make_descs(1, bg, Z, rm_length=E, rv=rv, mod_order=mod_order, Nref=Nref, nof_filler_bits=fillers)
        out = np.zeros(max(E, 1), np.uint8)
        self._check(self._L.ldpc_cuda_rate_match(self._h, _ptr(d), _ptr(codeblock), _ptr(out)), "ldpc_cuda_rate_match")
        return out[:E]

    def encode_batch_device(self, descs, msg_ptr, out_ptr):
        descs = np.ascontiguousarray(descs, CB_DESC)
        self._check(self._L.ldpc_cuda_encode_batch_device(self._h, descs.size, _ptr(descs), msg_ptr, out_ptr),
                    "ldpc_cuda_encode_batch_device")

    def rate_match_batch_device(self, descs, cw_ptr, out_ptr):
        descs = np.ascontiguousarray(descs, CB_DESC)
        self._check(self._L.ldpc_cuda_rate_match_batch_device(self._h, descs.size, _ptr(descs), cw_ptr, out_ptr),
                    "ldpc_cuda_rate_match_batch_device")

    # -- soft-buffer arena ---------------------------------------------------------------------------------------------
    def softbuffer_reserve(self, nbytes):
        self._check(self._L.ldpc_cuda_softbuffer_reserve(self._h, nbytes), "ldpc_cuda_softbuffer_reserve")

    def softbuffer_zero(self, offset, nbytes):
        self._check(self._L.ldpc_cuda_softbuffer_zero(self._h, offset, nbytes), "ldpc_cuda_softbuffer_zero")

    def softbuffer_read(self, offset, nbytes):
        dst = np.zeros(nbytes, np.int8)
        self._check(self._L.ldpc_cuda_softbuffer_read(self._h, offset, _ptr(dst), nbytes), "ldpc_cuda_softbuffer_read")
        return dst

    def softbuffer_write(self, offset, src):
        src = np.ascontiguousarray(src, np.int8)
        self._check(self._L.ldpc_cuda_softbuffer_write(self._h, offset, _ptr(src), src.size),
                    "ldpc_cuda_softbuffer_write")

    @property
    def softbuffer_ptr(self):
        return self._L.ldpc_cuda_softbuffer_ptr(self._h)


class LdpcCudaPool:
    """Several GPUs of one box from one process (ldpc_cuda_pool_*): one handle + one host thread per device, codeblocks of
    a decode-only batch round-robin (i -> device i mod G), transport blocks whole with HARQ affinity. Host buffers."""

    def __init__(self, devices, max_codeblocks_per_device=1024, flags=0, max_iterations=6, scaling_factor=0.8,
                 scale_mode=SCALE_GENERIC):
        self._L = library()
        devs = (ctypes.c_int * len(devices))(*devices)
        p = ctypes.c_void_p()
        rc = self._L.ldpc_cuda_pool_create(ctypes.byref(p), devs, len(devices), max_codeblocks_per_device, flags)
        if rc != 0:
            raise LdpcCudaError(rc, "ldpc_cuda_pool_create")
        self._p = p
        self.devices = list(devices)
        alg = _Algorithm(max_iterations, scaling_factor, scale_mode, 0)
        self._check(self._L.ldpc_cuda_pool_set_algorithm(self._p, ctypes.byref(alg)), "ldpc_cuda_pool_set_algorithm")

    def close(self):
        if getattr(self, "_p", None):
            self._L.ldpc_cuda_pool_destroy(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, where):
        if rc != 0:
            raise LdpcCudaError(rc, where)

    @property
    def nof_devices(self):
        return self._L.ldpc_cuda_pool_nof_devices(self._p)

    @property
    def launch_count(self):
        return self._L.ldpc_cuda_pool_launch_count(self._p)

    def codeblocks_per_device(self):
        out = np.zeros(self.nof_devices, np.uint64)
        self._check(self._L.ldpc_cuda_pool_codeblocks_per_device(self._p, _ptr(out), out.size),
                    "ldpc_cuda_pool_codeblocks_per_device")
        return out

    def device_of(self, harq_key):
        return self._L.ldpc_cuda_pool_device_of(self._p, harq_key)

    def softbuffer_reserve(self, nbytes_per_device):
        self._check(self._L.ldpc_cuda_pool_softbuffer_reserve(self._p, nbytes_per_device),
                    "ldpc_cuda_pool_softbuffer_reserve")

    def softbuffer_read(self, index, offset, nbytes):
        """Soft bits in the arena of device number `index`."""
        dst = np.zeros(nbytes, np.int8)
        h = self._L.ldpc_cuda_pool_handle(self._p, index)
        self._check(self._L.ldpc_cuda_softbuffer_read(h, offset, _ptr(dst), nbytes), "ldpc_cuda_softbuffer_read")
        return dst

    def decode_batch(self, descs, llrs, out, results=None):
        descs = np.ascontiguousarray(descs, CB_DESC)
        llrs = np.ascontiguousarray(llrs, np.int8)
        if results is None:
            results = np.zeros(descs.size, CB_RESULT)
        self._check(self._L.ldpc_cuda_pool_decode_batch(self._p, descs.size, _ptr(descs), _ptr(llrs), _ptr(out),
                                                        _ptr(results)), "ldpc_cuda_pool_decode_batch")
        return out, results

    def decode_batch_device(self, descs, llr_ptrs, out_ptrs, result_ptrs):
        """Inputs / outputs in device memory, one pointer per device (raw addresses); asynchronous: call synchronize()."""
        descs = np.ascontiguousarray(descs, CB_DESC)
        arr = lambda ptrs: (ctypes.c_void_p * len(ptrs))(*ptrs)
        a, b, c = arr(llr_ptrs), arr(out_ptrs), arr(result_ptrs)
        self._check(self._L.ldpc_cuda_pool_decode_batch_device(self._p, descs.size, _ptr(descs), a, b, c),
                    "ldpc_cuda_pool_decode_batch_device")

    def synchronize(self):
        self._check(self._L.ldpc_cuda_pool_synchronize(self._p), "ldpc_cuda_pool_synchronize")

    def dematch_decode_tb_batch(self, descs, rm_llrs, out, tbs, harq_keys, tb_out, results=None, tb_results=None):
        descs = np.ascontiguousarray(descs, CB_DESC)
        tbs = np.ascontiguousarray(tbs, TB_DESC)
        rm_llrs = np.ascontiguousarray(rm_llrs, np.int8)
        harq_keys = np.ascontiguousarray(harq_keys, np.uint64)
        assert harq_keys.size == tbs.size
        if results is None:
            results = np.zeros(descs.size, CB_RESULT)
        if tb_results is None:
            tb_results = np.zeros(tbs.size, TB_RESULT)
        self._check(self._L.ldpc_cuda_pool_dematch_decode_tb_batch(
            self._p, descs.size, _ptr(descs), _ptr(rm_llrs), None if out is None else _ptr(out), _ptr(results), tbs.size,
            _ptr(tbs), _ptr(harq_keys), _ptr(tb_out), _ptr(tb_results)), "ldpc_cuda_pool_dematch_decode_tb_batch")
        return out, results, tb_out, tb_results
