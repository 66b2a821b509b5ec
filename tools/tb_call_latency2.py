"""Experiment: what makes the 16-TB host call slow inside the aggregator? Variants: pinned by torch / by ldpc_cuda_host_alloc,
arena 1x (L2 resident) / 8x rotating."""
import ctypes, os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import __graft_entry__ as ge
import oracles as O
pkg = ge.load_package()
from importlib import import_module
L = import_module(pkg.__name__ + ".ldpc").library()
L.ldpc_cuda_host_alloc.argtypes = [ctypes.c_uint64, ctypes.c_uint32, ctypes.POINTER(ctypes.c_void_p)]

def host_alloc(a, flags=0):
    p = ctypes.c_void_p()
    assert L.ldpc_cuda_host_alloc(a.nbytes, flags, ctypes.byref(p)) == 0
    buf = (ctypes.c_uint8 * a.nbytes).from_address(p.value)
    out = np.frombuffer(buf, dtype=np.uint8).view(a.dtype).reshape(a.shape)
    out[...] = a
    return out

def torch_pin(a):
    t = torch.from_numpy(a.view(np.uint8) if a.dtype.fields else a).pin_memory()
    keep.append(t)
    r = t.numpy()
    return r.view(a.dtype) if a.dtype.fields else r
keep = []
rng = np.random.default_rng(3)
tbs, Qm, layers, nsym, esn0 = 1277992, 8, 4, 42588 * 4, 9.0
tb, cws, meta = O.build_transport_block(rng, tbs, 1, Qm, layers, nsym)
C = len(cws)
tx = np.concatenate([O.port().rate_match(cws[r], int(meta[r, 2]), 0, Qm, 0, int(meta[r, 3])) for r in range(C)])
llr1 = O.awgn_llrs(rng, tx, esn0)
d1, crc_bits = pkg.segment(tbs, 1, Qm, layers, nsym)
kb = 1056
ntb = 16
from concurrent.futures import ThreadPoolExecutor
ex = ThreadPoolExecutor(16)
for pin_name, pin in (("host_alloc", host_alloc),):
    for arena_mult, rewrite, wc in ((8, 16, 0), (8, -16, 0), (8, -1, 0)):
        dec = pkg.LdpcCuda(device=0, max_codeblocks=ntb * C)
        dec.set_algorithm(6, 0.8, O.SCALE_GENERIC)
        dec.softbuffer_reserve(arena_mult * ntb * C * 66 * 384)
        descs = np.zeros(ntb * C, pkg.CB_DESC)
        tbd = np.zeros(ntb, pkg.TB_DESC)
        tb_stride = ((tbs + 24 + 7) // 8 + 63) // 64 * 64
        for t in range(ntb):
            descs[t * C:(t + 1) * C] = d1
            for r in range(C):
                d = descs[t * C + r]
                d["early_stop"], d["new_data"] = 1, 1
                d["in_offset"] = int(d1[r]["in_offset"]) + t * llr1.size
                d["soft_offset"] = (t * C + r) * 66 * 384
                d["out_offset"] = (t * C + r) * kb
            tbd[t]["first_codeblock"], tbd[t]["nof_codeblocks"], tbd[t]["tbs_bits"] = t * C, C, tbs
            tbd[t]["nof_cb_crc_bits"], tbd[t]["tb_offset"] = int(crc_bits[0]), t * tb_stride
        llrs = host_alloc(np.tile(llr1, ntb), wc); cb_out = pin(np.zeros(ntb * C * kb, np.uint8)); tb_out = pin(np.zeros(ntb * tb_stride, np.uint8))
        res = pin(np.zeros(ntb * C, pkg.CB_RESULT)); tbr = pin(np.zeros(ntb, pkg.TB_RESULT))
        descs_p = pin(descs); tbd_p = pin(tbd)
        base = descs["soft_offset"].copy()
        src = np.tile(llr1, ntb)
        other = np.zeros_like(src)
        ts = []
        for it in range(24):
            descs_p["soft_offset"] = base + (it % arena_mult) * ntb * C * 66 * 384
            if rewrite == 1:
                llrs[:] = src
            elif rewrite < 0:  # the threads copy somewhere else
                n = llrs.size // -rewrite
                list(ex.map(lambda k: np.copyto(other[k * n:(k + 1) * n], src[k * n:(k + 1) * n]), range(-rewrite)))
            elif rewrite > 1:
                n = llrs.size // rewrite
                list(ex.map(lambda k: np.copyto(llrs[k * n:(k + 1) * n], src[k * n:(k + 1) * n]), range(rewrite)))
            t0 = time.perf_counter()
            dec.dematch_decode_tb_batch(descs_p, llrs, cb_out, tbd_p, tb_out, res, tbr)
            ts.append(time.perf_counter() - t0)
        ts = np.sort(np.array(ts[4:]))
        print(pin_name, "arena x%d rewrite by %d threads, write-combined %d" % (arena_mult, rewrite, wc), "median %.0f us min %.0f us" % (np.median(ts) * 1e6, ts[0] * 1e6), flush=True)
        dec.close()
