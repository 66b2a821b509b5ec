"""ctypes bindings to the two parity oracles (TEST INFRASTRUCTURE, never imported by the product package).

* ``port``  - oracle/liboracle_port.so, the plain-C restatement (always available once built).
* ``ref``   - oracle/_ref/libsrsran_ref.so, the reference's own code compiled from /root/reference (None if not built).

Also holds the synthetic test-signal helpers (random message -> encode -> BPSK/AWGN -> int8 LLR quantisation as
log_likelihood_ratio::quantize, lib/phy/upper/log_likelihood_ratio.cpp:87-96).
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

c_int, c_uint, c_float, c_void_p, c_char_p, c_double = (
    ctypes.c_int, ctypes.c_uint, ctypes.c_float, ctypes.c_void_p, ctypes.c_char_p, ctypes.c_double)

BG_K = {1: 22, 2: 10}
BG_N_SHORT = {1: 66, 2: 50}
BG_M = {1: 46, 2: 42}
LIFTING_SIZES = [2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 22, 24, 26, 28, 30, 32, 36, 40, 44, 48,
                 52, 56, 60, 64, 72, 80, 88, 96, 104, 112, 120, 128, 144, 160, 176, 192, 208, 224, 240, 256, 288, 320,
                 352, 384]
CRC_NONE, CRC16, CRC24A, CRC24B = 0, 1, 2, 3
CRC_BITS = {CRC16: 16, CRC24A: 24, CRC24B: 24}
SCALE_GENERIC, SCALE_AVX, SCALE_NEON = 0, 1, 2


def _ptr(a):
    return a.ctypes.data_as(c_void_p) if a is not None else None


def build_port():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "port"])


class PortOracle:
    def __init__(self):
        path = os.path.join(ORACLE_DIR, "liboracle_port.so")
        if not os.path.exists(path):
            build_port()
        L = self.lib = ctypes.CDLL(path)
        L.oracle_crc_packed.restype = c_uint
        L.oracle_crc_packed.argtypes = [c_int, c_void_p, c_uint]
        L.oracle_scale.argtypes = [c_int, c_float, c_int]
        L.oracle_decode.argtypes = [c_int, c_int, c_uint, c_int, c_uint, c_float, c_int, c_void_p, c_uint, c_void_p]
        L.oracle_rate_dematch.argtypes = [c_void_p, c_uint, c_void_p, c_uint, c_int, c_int, c_int, c_uint, c_uint]
        L.oracle_segment_rx.argtypes = [c_uint, c_int, c_int, c_uint, c_uint, c_void_p, c_uint]
        L.oracle_encode.argtypes = [c_int, c_int, c_void_p, c_void_p, c_uint]
        L.oracle_rate_match.argtypes = [c_void_p, c_uint, c_void_p, c_uint, c_int, c_int, c_uint, c_uint]
        L.oracle_hard_decision.argtypes = [c_void_p, c_uint, c_void_p]
        L.oracle_demodulate_soft.argtypes = [c_int, c_int, c_uint, c_void_p, c_void_p, c_void_p]
        L.oracle_descramble.argtypes = [c_uint, c_uint, c_void_p, c_void_p]
        L.oracle_gold_sequence.argtypes = [c_uint, c_uint, c_void_p]
        L.oracle_descramble_placeholders.argtypes = [c_uint, c_uint, c_void_p, c_void_p, c_uint, c_uint, c_uint, c_void_p]
        L.oracle_ulsch_demultiplex.argtypes = [c_void_p, c_uint, c_uint, c_uint, c_void_p, c_uint, c_void_p, c_void_p,
                                               c_void_p, c_void_p, c_void_p, c_void_p]
        L.oracle_ulsch_placeholders.argtypes = [c_void_p, c_uint, c_uint, c_uint, c_uint, c_uint, c_uint, c_void_p, c_uint]

    def demodulate_soft(self, mod_order, symbols, noise_vars, pi2=False):
        """symbols: complex64 [n]; noise_vars: float32 [n]. Returns int8 [n * mod_order]."""
        symbols = np.ascontiguousarray(symbols, np.complex64)
        noise_vars = np.ascontiguousarray(noise_vars, np.float32)
        out = np.zeros(symbols.size * mod_order, np.int8)
        self.lib.oracle_demodulate_soft(mod_order, int(pi2), symbols.size, _ptr(symbols), _ptr(noise_vars), _ptr(out))
        return out

    def descramble(self, c_init, llrs):
        llrs = np.ascontiguousarray(llrs, np.int8)
        out = np.zeros_like(llrs)
        self.lib.oracle_descramble(c_init, llrs.size, _ptr(llrs), _ptr(out))
        return out

    def descramble_placeholders(self, c_init, llrs, mod_order, nof_layers, re_indexes):
        """pusch_demodulator_impl::descramble with the UCI repetition placeholders of the listed REs (ascending)."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        re_indexes = np.ascontiguousarray(re_indexes, np.uint16)
        out = np.zeros_like(llrs)
        self.lib.oracle_descramble_placeholders(c_init, llrs.size, _ptr(llrs), _ptr(out), mod_order, nof_layers,
                                                re_indexes.size, _ptr(re_indexes))
        return out

    @staticmethod
    def _ulsch_cfg(cfg):
        """cfg: dict(mod_order, nof_layers, nof_prb, start_symbol, nof_symbols, dmrs_type, cdm, dmrs_mask, rvd)."""
        return np.array([cfg["mod_order"], cfg["nof_layers"], cfg["nof_prb"], cfg["start_symbol"], cfg["nof_symbols"],
                         cfg["dmrs_type"], cfg["cdm"], cfg["dmrs_mask"], cfg["rvd"]], np.uint32)

    def ulsch_demultiplex(self, cfg, G_ack, G_csi1, G_csi2, llrs):
        """Returns (sch, ack, csi1, csi2) or None where the reference asserts."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        c = self._ulsch_cfg(cfg)
        sch = np.zeros(llrs.size + 64, np.int8)
        ack, c1, c2 = np.zeros(G_ack + 64, np.int8), np.zeros(G_csi1 + 64, np.int8), np.zeros(G_csi2 + 64, np.int8)
        used, slen = ctypes.c_uint(0), ctypes.c_uint(0)
        rc = self.lib.oracle_ulsch_demultiplex(_ptr(c), G_ack, G_csi1, G_csi2, _ptr(llrs), llrs.size, _ptr(sch), _ptr(ack),
                                               _ptr(c1), _ptr(c2), ctypes.byref(used), ctypes.byref(slen))
        if rc != 0 or used.value != llrs.size:
            return None
        return sch[:slen.value], ack[:G_ack], c1[:G_csi1], c2[:G_csi2]

    def ulsch_placeholders(self, cfg, O_ack, G_ack, O_csi1, G_csi1, O_csi2, G_csi2):
        c = self._ulsch_cfg(cfg)
        lst = np.zeros(1 << 16, np.uint16)
        n = self.lib.oracle_ulsch_placeholders(_ptr(c), O_ack, G_ack, O_csi1, G_csi1, O_csi2, G_csi2, _ptr(lst), lst.size)
        return None if n < 0 else lst[:n].copy()

    def gold_sequence(self, c_init, n):
        c = np.zeros(n, np.uint8)
        self.lib.oracle_gold_sequence(c_init, n, _ptr(c))
        return c

    def llr_add(self, a, b):
        return self.lib.oracle_llr_add(int(a), int(b))

    def llr_sub(self, a, b):
        return self.lib.oracle_llr_sub(int(a), int(b))

    def llr_promotion_sum(self, a, b):
        return self.lib.oracle_llr_promotion_sum(int(a), int(b))

    def hard_decision(self, soft):
        soft = np.ascontiguousarray(soft, np.int8)
        out = np.zeros((soft.size + 7) // 8, np.uint8)
        self.lib.oracle_hard_decision(_ptr(soft), soft.size, _ptr(out))
        return out

    def crc(self, poly, packed, nbits):
        packed = np.ascontiguousarray(packed, np.uint8)
        return self.lib.oracle_crc_packed(poly, _ptr(packed), nbits)

    def scale(self, m, sf=0.8, mode=SCALE_GENERIC):
        return self.lib.oracle_scale(int(m), sf, mode)

    def decode(self, bg, Z, llrs, fillers=0, crc=CRC_NONE, max_iters=6, sf=0.8, scale_mode=SCALE_GENERIC, out_init=0):
        llrs = np.ascontiguousarray(llrs, np.int8)
        out = np.full((BG_K[bg] * Z + 7) // 8, out_init, np.uint8)
        it = self.lib.oracle_decode(bg, Z, fillers, crc, max_iters, sf, scale_mode, _ptr(llrs), llrs.size, _ptr(out))
        return out, it

    def rate_dematch(self, soft, llrs, new_data, rv=0, Qm=1, Nref=0, fillers=0):
        """soft is modified in place (int8 array of N_short*Z)."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        assert soft.dtype == np.int8 and soft.flags.c_contiguous
        r = self.lib.oracle_rate_dematch(_ptr(soft), soft.size, _ptr(llrs), llrs.size, int(new_data), rv, Qm, Nref,
                                         fillers)
        assert r == 0, r
        return soft

    def segment_rx(self, tbs, bg, Qm, nof_layers, nof_ch_symbols):
        C = self.lib.oracle_segment_rx(tbs, bg, Qm, nof_layers, nof_ch_symbols, None, 0)
        assert C > 0, C
        meta = np.zeros((C, 6), np.uint32)
        self.lib.oracle_segment_rx(tbs, bg, Qm, nof_layers, nof_ch_symbols, _ptr(meta), C)
        return meta

    def encode(self, bg, Z, msg_bits, out_len=None):
        msg_bits = np.ascontiguousarray(msg_bits, np.uint8)
        assert msg_bits.size == BG_K[bg] * Z
        out_len = BG_N_SHORT[bg] * Z if out_len is None else out_len
        out = np.zeros(out_len, np.uint8)
        r = self.lib.oracle_encode(bg, Z, _ptr(msg_bits), _ptr(out), out_len)
        assert r == 0, r
        return out

    def rate_match(self, cw, E, rv=0, Qm=1, Nref=0, fillers=0):
        cw = np.ascontiguousarray(cw, np.uint8)
        out = np.zeros(E, np.uint8)
        r = self.lib.oracle_rate_match(_ptr(out), E, _ptr(cw), cw.size, rv, Qm, Nref, fillers)
        assert r == 0, r
        return out


class RefOracle:
    """The reference's own compiled code. One decoder / dematcher object per (type) is cached unless fresh=True."""

    def __init__(self, path):
        L = self.lib = ctypes.CDLL(path)
        L.ref_cpu_supports.argtypes = [c_char_p]
        L.ref_graph_dump.argtypes = [c_int, c_int, c_void_p, c_void_p]
        L.ref_llr_quantize.argtypes = [c_float, c_float]
        L.ref_hard_decision.argtypes = [c_void_p, c_uint, c_void_p]
        L.ref_crc_packed.restype = c_uint
        L.ref_crc_packed.argtypes = [c_char_p, c_int, c_void_p, c_uint]
        L.ref_decoder_create.restype = c_void_p
        L.ref_decoder_create.argtypes = [c_char_p]
        L.ref_decoder_destroy.argtypes = [c_void_p]
        L.ref_decoder_decode.argtypes = [c_void_p, c_int, c_int, c_uint, c_int, c_uint, c_float, c_void_p, c_uint,
                                         c_void_p]
        L.ref_decoder_decode_batch_mt.restype = c_double
        L.ref_decoder_decode_batch_mt.argtypes = [c_char_p, c_uint, c_uint, c_int, c_int, c_uint, c_int, c_uint,
                                                  c_float, c_void_p, c_uint, c_uint, c_void_p, c_uint, c_void_p]
        L.ref_dematch_decode_batch_mt.restype = c_double
        L.ref_dematch_decode_batch_mt.argtypes = [c_char_p, c_char_p, c_uint, c_uint, c_int, c_int, c_int, c_int, c_uint,
                                                  c_int, c_uint, c_void_p, c_uint, c_uint, c_void_p, c_uint, c_void_p]
        L.ref_dematcher_create.restype = c_void_p
        L.ref_dematcher_create.argtypes = [c_char_p]
        L.ref_dematcher_destroy.argtypes = [c_void_p]
        L.ref_dematcher_run.argtypes = [c_void_p, c_void_p, c_uint, c_void_p, c_uint, c_int, c_int, c_int, c_int,
                                        c_int, c_uint, c_uint]
        L.ref_encode.argtypes = [c_char_p, c_int, c_int, c_void_p, c_void_p, c_uint]
        L.ref_rate_match.argtypes = [c_void_p, c_uint, c_void_p, c_uint, c_int, c_int, c_int, c_int, c_uint, c_uint]
        L.ref_segment_rx.argtypes = [c_uint, c_int, c_int, c_int, c_uint, c_uint, c_uint, c_void_p]
        L.ref_segment_tx.argtypes = [c_void_p, c_uint, c_int, c_int, c_int, c_uint, c_uint, c_uint, c_void_p, c_uint]
        L.ref_pusch_create.restype = c_void_p
        L.ref_pusch_create.argtypes = [c_char_p, c_char_p, c_uint, c_uint, c_uint]
        L.ref_pusch_destroy.argtypes = [c_void_p]
        L.ref_pusch_decode.argtypes = [c_void_p, c_uint, c_uint, c_uint, c_void_p, c_uint, c_void_p, c_uint, c_int,
                                       c_int, c_int, c_uint, c_uint, c_uint, c_int, c_int, c_int, c_void_p]
        L.ref_pusch_concurrent.restype = c_double
        L.ref_pusch_concurrent.argtypes = [c_char_p, c_char_p, c_uint, c_uint, c_uint, c_uint, c_void_p, c_uint, c_int, c_int,
                                           c_uint, c_uint, c_int, c_uint, c_uint, c_void_p, c_void_p, c_void_p]
        L.ref_demodulate_soft.argtypes = [c_int, c_int, c_uint, c_void_p, c_void_p, c_void_p, c_uint]
        L.ref_ulsch_demultiplex.argtypes = [c_void_p, c_void_p, c_uint, c_void_p, c_uint, c_void_p, c_uint, c_void_p, c_uint,
                                            c_void_p, c_uint]
        L.ref_ulsch_placeholders.argtypes = [c_void_p, c_void_p, c_void_p, c_uint]
        L.ref_descramble.argtypes = [c_uint, c_uint, c_void_p, c_void_p]
        L.ref_descramble_placeholders.argtypes = [c_uint, c_uint, c_uint, c_void_p, c_void_p, c_uint, c_uint, c_uint, c_void_p]
        if hasattr(L, "dropin_ulsch_demultiplex"):  # oracle/_ref/libdropin.so only
            L.dropin_ulsch_demultiplex.argtypes = L.ref_ulsch_demultiplex.argtypes
            L.dropin_ulsch_placeholders.argtypes = L.ref_ulsch_placeholders.argtypes
        if hasattr(L, "dropin_descramble_placeholders"):  # oracle/_ref/libdropin.so only
            L.dropin_descramble_placeholders.argtypes = [c_uint, c_uint, c_uint, c_void_p, c_void_p, c_uint, c_uint, c_uint,
                                                         c_void_p]
        self._decoders = {}
        self._dematchers = {}

    def pusch_concurrent(self, decoder_type, pool_type, llrs, tb_bytes, bg, Qm, nof_layers, max_iters=6, early_stop=True,
                         aggregate_blocks=3, warm_reps=0):
        """K pusch_decoder objects of one factory on K threads sharing one soft-buffer pool (uplink_processor_impl.cpp:
        155-171). llrs: int8 [K][reps][nof_tx][nof_llrs]. Returns (seconds, tb_out [K][reps][tb_bytes], stats
        [K][reps][nof_tx][6], (launches, transport blocks, max per launch, then join-wait / stage / follow / drain / launch ns) of the
        product's slot aggregator)."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        K, reps, nof_tx, nof_llrs = llrs.shape
        tb_out = np.zeros((K, reps, tb_bytes), np.uint8)
        stats = np.zeros((K, reps, nof_tx, 6), np.uint32)
        agg = np.zeros(8, np.uint64)
        secs = self.lib.ref_pusch_concurrent(decoder_type.encode(), pool_type.encode(), K, reps, nof_tx, tb_bytes, _ptr(llrs),
                                             nof_llrs, bg, Qm, nof_layers, max_iters, int(early_stop), aggregate_blocks,
                                             warm_reps, _ptr(tb_out), _ptr(stats), _ptr(agg))
        assert secs >= 0, "ref_pusch_concurrent(%s, %s) failed: %g" % (decoder_type, pool_type, secs)
        return secs, tb_out, stats, tuple(int(x) for x in agg)

    def demodulate_soft(self, mod_order, symbols, noise_vars, pi2=False, batch=0):
        """demodulation_mapper_impl::demodulate_soft; batch > 0 feeds `batch` symbols per call (below the AVX2 width: the
        reference's generic per-symbol code only)."""
        symbols = np.ascontiguousarray(symbols, np.complex64)
        noise_vars = np.ascontiguousarray(noise_vars, np.float32)
        out = np.zeros(symbols.size * mod_order, np.int8)
        rc = self.lib.ref_demodulate_soft(mod_order, int(pi2), symbols.size, _ptr(symbols), _ptr(noise_vars), _ptr(out),
                                          batch)
        assert rc == 0
        return out

    def descramble(self, c_init, llrs):
        llrs = np.ascontiguousarray(llrs, np.int8)
        out = np.zeros_like(llrs)
        self.lib.ref_descramble(c_init, llrs.size, _ptr(llrs), _ptr(out))
        return out

    def descramble_placeholders(self, rnti, n_id, llrs, mod_order, nof_layers, re_indexes, cuda=False):
        """The reference's own pusch_demodulator_impl::descramble (c_init = rnti * 2^15 + n_id); cuda=True (drop-in library):
        the product's descramble_cuda(mapper, out, in, config) adaptor on the same configuration."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        re_indexes = np.ascontiguousarray(re_indexes, np.uint16)
        out = np.zeros_like(llrs)
        if cuda:
            rc = self.lib.dropin_descramble_placeholders(rnti, n_id, llrs.size, _ptr(llrs), _ptr(out), mod_order, nof_layers,
                                                         re_indexes.size, _ptr(re_indexes))
            assert rc == 0
            return out
        self.lib.ref_descramble_placeholders(rnti, n_id, llrs.size, _ptr(llrs), _ptr(out), mod_order, nof_layers,
                                             re_indexes.size, _ptr(re_indexes))
        return out

    def ulsch_demultiplex(self, cfg, llrs, sch_len, G_ack, G_csi1, G_csi2, cuda=False):
        """The reference's own ulsch_demultiplex_impl::demultiplex (sizes must be consistent: it asserts otherwise);
        cuda=True (drop-in library): the product's create_ulsch_demultiplex_cuda() behind the same interface."""
        llrs = np.ascontiguousarray(llrs, np.int8)
        c = PortOracle._ulsch_cfg(cfg)
        sch, ack = np.zeros(sch_len, np.int8), np.zeros(G_ack, np.int8)
        c1, c2 = np.zeros(G_csi1, np.int8), np.zeros(G_csi2, np.int8)
        if cuda:
            rc = self.lib.dropin_ulsch_demultiplex(_ptr(c), _ptr(llrs), llrs.size, _ptr(sch), sch_len, _ptr(ack), G_ack,
                                                   _ptr(c1), G_csi1, _ptr(c2), G_csi2)
            assert rc == 0
            return sch, ack, c1, c2
        self.lib.ref_ulsch_demultiplex(_ptr(c), _ptr(llrs), llrs.size, _ptr(sch), sch_len, _ptr(ack), G_ack, _ptr(c1), G_csi1,
                                       _ptr(c2), G_csi2)
        return sch, ack, c1, c2

    def ulsch_placeholders(self, cfg, O_ack, G_ack, O_csi1, G_csi1, O_csi2, G_csi2, cuda=False):
        c = PortOracle._ulsch_cfg(cfg)
        info = np.array([O_ack, G_ack, O_csi1, G_csi1, O_csi2, G_csi2], np.uint32)
        lst = np.zeros(1 << 16, np.uint16)
        fn = self.lib.dropin_ulsch_placeholders if cuda else self.lib.ref_ulsch_placeholders
        n = fn(_ptr(c), _ptr(info), _ptr(lst), lst.size)
        assert n >= 0
        return lst[:n].copy()

    def supports(self, kind):
        return bool(self.lib.ref_cpu_supports(kind.encode()))

    def llr_add(self, a, b):
        return self.lib.ref_llr_add(int(a), int(b))

    def llr_sub(self, a, b):
        return self.lib.ref_llr_sub(int(a), int(b))

    def llr_promotion_sum(self, a, b):
        return self.lib.ref_llr_promotion_sum(int(a), int(b))

    def quantize(self, v, rng):
        return self.lib.ref_llr_quantize(float(v), float(rng))

    def hard_decision(self, soft):
        soft = np.ascontiguousarray(soft, np.int8)
        out = np.zeros((soft.size + 7) // 8, np.uint8)
        self.lib.ref_hard_decision(_ptr(soft), soft.size, _ptr(out))
        return out

    def crc(self, poly, packed, nbits, impl="auto"):
        packed = np.ascontiguousarray(packed, np.uint8)
        return self.lib.ref_crc_packed(impl.encode(), poly, _ptr(packed), nbits)

    def graph(self, bg, Z):
        M, N = BG_M[bg], BG_K[bg] + BG_M[bg]
        pcm = np.zeros((M, N), np.uint16)
        adj = np.zeros((M, 20), np.uint16)
        self.lib.ref_graph_dump(bg, Z, _ptr(pcm), _ptr(adj))
        return pcm, adj

    def decode(self, bg, Z, llrs, fillers=0, crc=CRC_NONE, max_iters=6, sf=0.8, kind="generic", fresh=True,
               out_init=0):
        llrs = np.ascontiguousarray(llrs, np.int8)
        out = np.full((BG_K[bg] * Z + 7) // 8, out_init, np.uint8)
        if fresh:
            h = self.lib.ref_decoder_create(kind.encode())
        else:
            h = self._decoders.get(kind) or self._decoders.setdefault(kind, self.lib.ref_decoder_create(kind.encode()))
        assert h, "reference decoder type %s unavailable" % kind
        it = self.lib.ref_decoder_decode(h, bg, Z, fillers, crc, max_iters, sf, _ptr(llrs), llrs.size, _ptr(out))
        if fresh:
            self.lib.ref_decoder_destroy(h)
        return out, it

    def decode_batch_mt(self, kind, nthreads, bg, Z, llrs2d, fillers=0, crc=CRC_NONE, max_iters=6, sf=0.8):
        llrs2d = np.ascontiguousarray(llrs2d, np.int8)
        ncb, n = llrs2d.shape
        kb = (BG_K[bg] * Z + 7) // 8
        out = np.zeros((ncb, kb), np.uint8)
        iters = np.zeros(ncb, np.int32)
        secs = self.lib.ref_decoder_decode_batch_mt(kind.encode(), nthreads, ncb, bg, Z, fillers, crc, max_iters, sf,
                                                    _ptr(llrs2d), n, n, _ptr(out), kb, _ptr(iters))
        assert secs >= 0, "reference decoder type %s unavailable" % kind
        return secs, out, iters

    def dematch_decode_batch_mt(self, kind, nthreads, bg, Z, rm2d, rv=0, Qm=1, fillers=0, crc=CRC_NONE, max_iters=6):
        """rate_dematch (new data) + decode with CRC early stop per codeblock, nthreads workers on disjoint slices (what
        pusch_decoder_impl::decode does per codeblock, pusch_decoder_impl.cpp:156-201). rm2d: [ncb, E] int8."""
        rm2d = np.ascontiguousarray(rm2d, np.int8)
        ncb, E = rm2d.shape
        kb = (BG_K[bg] * Z + 7) // 8
        out = np.zeros((ncb, kb), np.uint8)
        iters = np.zeros(ncb, np.int32)
        dm = kind if kind in ("avx2", "avx512") else "generic"
        secs = self.lib.ref_dematch_decode_batch_mt(dm.encode(), kind.encode(), nthreads, ncb, bg, Z, rv, Qm, fillers, crc,
                                                    max_iters, _ptr(rm2d), E, E, _ptr(out), kb, _ptr(iters))
        assert secs >= 0, "reference type %s unavailable" % kind
        return secs, out, iters

    def rate_dematch(self, soft, llrs, new_data, bg, Z, rv=0, Qm=1, Nref=0, fillers=0, kind="generic"):
        llrs = np.ascontiguousarray(llrs, np.int8)
        assert soft.dtype == np.int8 and soft.flags.c_contiguous
        h = self._dematchers.get(kind) or self._dematchers.setdefault(kind,
                                                                      self.lib.ref_dematcher_create(kind.encode()))
        assert h
        self.lib.ref_dematcher_run(h, _ptr(soft), soft.size, _ptr(llrs), llrs.size, int(new_data), bg, Z, rv, Qm, Nref,
                                   fillers)
        return soft

    def encode(self, bg, Z, msg_bits, out_len=None, kind="auto"):
        msg_bits = np.ascontiguousarray(msg_bits, np.uint8)
        out_len = BG_N_SHORT[bg] * Z if out_len is None else out_len
        out = np.zeros(out_len, np.uint8)
        r = self.lib.ref_encode(kind.encode(), bg, Z, _ptr(msg_bits), _ptr(out), out_len)
        assert r == 0
        return out

    def rate_match(self, cw, E, bg, Z, rv=0, Qm=1, Nref=0, fillers=0):
        cw = np.ascontiguousarray(cw, np.uint8)
        out = np.zeros(E, np.uint8)
        self.lib.ref_rate_match(_ptr(out), E, _ptr(cw), cw.size, bg, Z, rv, Qm, Nref, fillers)
        return out

    def segment_rx(self, tbs, bg, Qm, nof_layers, nof_ch_symbols, rv=0, Nref=0):
        meta = np.zeros((52, 6), np.uint32)
        C = self.lib.ref_segment_rx(tbs, bg, rv, Qm, Nref, nof_layers, nof_ch_symbols, _ptr(meta))
        return None if C < 0 else meta[:C].copy()

    def segment_tx(self, tb_bytes, bg, Qm, nof_layers, nof_ch_symbols, rv=0, Nref=0):
        tb_bytes = np.ascontiguousarray(tb_bytes, np.uint8)
        segs = np.zeros((52, 8448), np.uint8)
        C = self.lib.ref_segment_tx(_ptr(tb_bytes), tb_bytes.size, bg, rv, Qm, Nref, nof_layers, nof_ch_symbols,
                                    _ptr(segs), 8448)
        return None if C < 0 else segs[:C].copy()


_port = None
_ref = None
_ref_tried = False


def port():
    global _port
    if _port is None:
        _port = PortOracle()
    return _port


def ref():
    """Returns the compiled-reference oracle or None when oracle/_ref was never built (no /root/reference)."""
    global _ref, _ref_tried
    if not _ref_tried:
        _ref_tried = True
        path = os.path.join(ORACLE_DIR, "_ref", "libsrsran_ref.so")
        if os.path.exists(path):
            try:
                _ref = RefOracle(path)
            except OSError:
                _ref = None
    return _ref


# ----------------------------------------------------------------------------------------------------------------------
# Synthetic signals.
def quantize_llr(x, range_limit):
    """log_likelihood_ratio::quantize (log_likelihood_ratio.cpp:87-96): clip, scale to +-120, std::round (half away)."""
    x = np.clip(np.asarray(x, np.float32), -range_limit, range_limit) / np.float32(range_limit) * np.float32(120)
    return (np.sign(x) * np.floor(np.abs(x) + np.float32(0.5))).astype(np.int8)


def random_message(rng, bg, Z, fillers=0):
    K = BG_K[bg] * Z
    msg = rng.integers(0, 2, K).astype(np.uint8)
    if fillers:
        msg[K - fillers:] = 254
    return msg


def attach_crc(msg, poly, fillers=0):
    """Overwrites the last CRC_BITS[poly] payload bits (before the fillers) with the CRC of what precedes them."""
    nb = CRC_BITS[poly]
    K = msg.size - fillers
    payload = msg[:K - nb]
    packed = np.packbits(payload)
    c = port().crc(poly, packed, payload.size)
    msg[K - nb:K] = [(c >> (nb - 1 - i)) & 1 for i in range(nb)]
    return msg


def awgn_llrs(rng, code_bits, esn0_db, range_limit=20.0):
    """BPSK (0 -> +1) over AWGN, LLR = 2y/sigma^2 quantised to int8 as SURVEY.md 8(d) cfg 1. Filler (254) -> +127."""
    code_bits = np.asarray(code_bits)
    sigma2 = 10.0 ** (-esn0_db / 10.0) / 2.0
    s = 1.0 - 2.0 * (code_bits & 1).astype(np.float32)
    y = s + rng.normal(0.0, np.sqrt(sigma2), code_bits.shape).astype(np.float32)
    llr = quantize_llr(2.0 * y / sigma2, range_limit)
    llr[code_bits == 254] = 127
    return llr


def make_codeword(rng, bg, Z, fillers=0, crc=CRC_NONE):
    msg = random_message(rng, bg, Z, fillers)
    if crc != CRC_NONE:
        attach_crc(msg, crc, fillers)
    cw = port().encode(bg, Z, msg)
    return msg, cw


MOD_NORM = {1: 2.0, 2: 2.0, 4: 10.0, 6: 42.0, 8: 170.0}


def random_symbols(rng, mod_order, n, noise_var=0.01, spread=1.3):
    """Equalised symbols for demodulator tests: constellation points (not bit-mapped: any grid point) + complex noise, a
    few outliers far outside the constellation, and per-symbol noise variances over three decades (plus some <= 0, for
    which the reference emits 0)."""
    levels = {1: 1, 2: 1, 4: 3, 6: 7, 8: 15}[mod_order]
    pts = np.arange(-levels, levels + 1, 2)
    d = 1.0 / np.sqrt(MOD_NORM[mod_order])
    re, im = rng.choice(pts, n) * d, rng.choice(pts, n) * d
    sym = (re + 1j * im + (rng.normal(0, 1, n) + 1j * rng.normal(0, 1, n)) * np.sqrt(noise_var / 2)).astype(np.complex64)
    far = rng.random(n) < 0.01
    sym[far] *= np.float32(spread * 2)
    nv = (noise_var * 10.0 ** rng.uniform(-1.5, 1.5, n)).astype(np.float32)
    nv[rng.random(n) < 0.005] = 0.0
    nv[rng.random(n) < 0.002] = -1.0
    return sym, nv


def msg_to_packed(msg):
    return np.packbits(np.where(msg == 254, 0, msg).astype(np.uint8))


def build_transport_block(rng, tbs_bits, bg, Qm, nof_layers, nof_ch_symbols):
    """TS 38.212 5.2.2 segmentation of a random transport block with several codeblocks, done with the oracle port (the
    reference's TX segmenter stops at 52 codeblocks): TB + CRC24A split evenly, CRC24B per codeblock, fillers, encode.
    Returns (tb bytes, list of codewords, meta[C, 6] from segment_rx)."""
    meta = port().segment_rx(tbs_bits, bg, Qm, nof_layers, nof_ch_symbols)
    C = len(meta)
    assert C > 1 and tbs_bits % 8 == 0
    tb_bits = rng.integers(0, 2, tbs_bits).astype(np.uint8)
    crc = port().crc(CRC24A, np.packbits(tb_bits), tbs_bits)
    tb_all = np.concatenate([tb_bits, np.array([(crc >> (23 - k)) & 1 for k in range(24)], np.uint8)])
    Z, F = int(meta[0, 0]), int(meta[0, 3])
    K = BG_K[bg] * Z
    per_cb = K - F - 24
    assert per_cb * C == tb_all.size, (per_cb, C, tb_all.size)
    cws = []
    for r in range(C):
        msg = np.concatenate([tb_all[r * per_cb:(r + 1) * per_cb], np.zeros(24, np.uint8), np.full(F, 254, np.uint8)])
        cws.append(port().encode(bg, Z, attach_crc(msg, CRC24B, F)))
    return np.packbits(tb_bits), cws, meta


def ulsch_cases(rng, n):
    """Random UL-SCH multiplexing configurations over the space ulsch_demultiplex takes (TS 38.212 6.2.7): modulation, 1-4
    layers, DM-RS type and positions, HARQ-ACK with / without reserved elements (O_ACK <= 2 reserves), CSI part 1 / 2.
    Yields (cfg, O_ack, G_ack, O_csi1, G_csi1, O_csi2, G_csi2); inconsistent draws are left to the caller to skip."""
    masks = [0b0000000000000100, 0b0000100000000100, 0b0000000000001100, 0b0000100100000100, 0b0000000000000001]
    out = []
    while len(out) < n:
        Qm = int(rng.choice([1, 2, 4, 6, 8]))
        nl = int(rng.integers(1, 5))
        bpr = Qm * nl
        dmrs_type = int(rng.integers(1, 3))
        cfg = dict(mod_order=Qm, nof_layers=nl, nof_prb=int(rng.choice([1, 2, 5, 24, 52, 106, 273])),
                   start_symbol=int(rng.choice([0, 0, 1, 2])), nof_symbols=0, dmrs_type=dmrs_type,
                   cdm=int(rng.integers(1, 3 if dmrs_type == 1 else 4)), dmrs_mask=int(rng.choice(masks)), rvd=0)
        cfg["nof_symbols"] = int(rng.integers(4, 15 - cfg["start_symbol"]))
        re_total = cfg["nof_prb"] * 12 * cfg["nof_symbols"]
        mode = int(rng.integers(0, 4))
        O_ack = int(rng.choice([0, 1, 2, 3, 11])) if mode else 0
        G_ack = bpr * int(rng.integers(1, max(2, re_total // 20))) if O_ack else 0
        if O_ack in (1, 2) or (O_ack == 0 and rng.integers(0, 3) == 0):
            cfg["rvd"] = max(G_ack, bpr * int(rng.integers(1, max(2, re_total // 16))))
        O_c1 = int(rng.choice([0, 1, 4, 20]))
        G_c1 = bpr * int(rng.integers(1, max(2, re_total // 12))) if O_c1 else 0
        O_c2 = int(rng.choice([0, 1, 7])) if O_c1 else 0
        G_c2 = bpr * int(rng.integers(1, max(2, re_total // 12))) if O_c2 else 0
        out.append((cfg, O_ack, G_ack, O_c1, G_c1, O_c2, G_c2))
    return out
