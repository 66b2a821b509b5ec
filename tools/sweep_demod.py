"""Development helper: demodulator + descrambler throughput per modulation order on one B200 (device-resident inputs,
about 174 M LLRs per launch as in bench.py's `demodulator` object). Prints a markdown table."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge

pkg = ge.load_package()
dev = torch.device("cuda", 0)
dec = pkg.LdpcCuda(device=0, max_codeblocks=64)
stream = torch.cuda.ExternalStream(dec.stream, device=dev)  # the events must sit on the stream the kernels are launched on
n_jobs, n_llr_job = 128, 1362816
print("| modulation | LLRs / launch | ms | G LLR/s | algorithmic GB/s |")
print("|---|---|---|---|---|")
for q, name in [(8, "256QAM"), (6, "64QAM"), (4, "16QAM"), (2, "QPSK"), (1, "BPSK"), (0, "descramble only")]:
    n_sym = n_llr_job // max(q, 1)
    g = torch.Generator(device=dev)
    g.manual_seed(5)
    sym = torch.randn((n_jobs * n_sym, 2), generator=g, device=dev).contiguous() if q else None
    nv = torch.full((n_jobs * n_sym,), 0.05, dtype=torch.float32, device=dev) if q else None
    in_llrs = torch.randint(-120, 121, (n_jobs * n_sym,), generator=g, device=dev, dtype=torch.int8) if not q else None
    out = torch.empty(n_jobs * n_sym * max(q, 1), dtype=torch.int8, device=dev)
    jobs = np.zeros(n_jobs, pkg.DEMOD_DESC)
    jobs["mod_order"], jobs["descramble"], jobs["nof_symbols"] = q, 1, n_sym
    jobs["c_init"] = (np.arange(n_jobs) + 1) * 32768 + 17
    jobs["symbol_offset"] = np.arange(n_jobs, dtype=np.uint64) * np.uint64(n_sym)
    jobs["llr_offset"] = np.arange(n_jobs, dtype=np.uint64) * np.uint64(n_sym * max(q, 1))
    args = (jobs, sym.data_ptr() if q else 0, nv.data_ptr() if q else 0, in_llrs.data_ptr() if not q else 0, out.data_ptr())
    torch.cuda.synchronize()
    for _ in range(3):
        dec.demodulate_batch_device(*args)
    dec.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(5):
        dec.demodulate_batch_device(*args)
    e1.record(stream)
    dec.synchronize()
    ms = e0.elapsed_time(e1) / 5
    n = n_jobs * n_sym * max(q, 1)
    by = n_jobs * n_sym * ((12 + q) if q else 2)
    print("| %s | %d | %.3f | %.0f | %.0f |" % (name, n, ms, n / ms / 1e6, by / ms / 1e6))
dec.close()
