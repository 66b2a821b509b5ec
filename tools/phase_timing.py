"""Where a codeblock pair's time goes inside the decoder kernel (development aid). Needs a library built with
-DTM_PHASE_TIMING:
    tools/build_variant.sh srsran_project_23.5_b200/csrc/ldpc_kernels_tmem.cuh build/phase.so -DTM_PHASE_TIMING
    SRSRAN_LDPC_CUDA_LIB=$PWD/build/phase.so python tools/phase_timing.py
Runs bench.py's fused full-slot workload (pusch_slot) and the plain decode workload on a small batch and prints the share
of cycles thread 0 of every block spent per phase."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

NAMES = ["claim + work records", "fused: stage in", "fused: scatter / combine", "load / stage out + convert",
         "input size, TMEM reset, prefetch", "row passes", "check: sweep over the hard bits", "pair end barrier",
         "check: reduction + barrier", "finish: message bytes", "finish: output + barrier",
         "fused: scatter A (thread 0)", "fused: scatter B (thread 0)"]


def main():
    import torch
    pkg = ge.load_package()
    from importlib import import_module
    L = import_module(pkg.__name__ + ".ldpc").library()
    out = (ctypes.c_ulonglong * 16)()
    import bench
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    dec = pkg.LdpcCuda(device=0, max_codeblocks=1 << 17, max_iterations=6)
    stream = torch.cuda.ExternalStream(dec.stream, device=dev)

    def report(title):
        torch.cuda.synchronize()
        assert L.ldpc_cuda_debug_phase_cycles(out) == 0
        total = float(sum(out[:13])) or 1.0
        print(title)
        for i, nm in enumerate(NAMES):
            print("  %-36s %6.2f %%" % (nm, 100.0 * out[i] / total))

    # fused full-slot workload (BASELINE configs 3 + 5)
    assert L.ldpc_cuda_debug_phase_cycles(out) == 0
    r = bench.pusch_slot_extra(torch, pkg, dec, dev, stream, 2)
    report("pusch_slot (fused, R = 0.94, device resident): %.1f Gb/s" % r["value"])
    # plain decode, fixed 6 iterations and early stop
    llr_d = bench.make_workload_gpu(torch, pkg, dec, dev, 16384, -1.0, 1)
    n = llr_d.shape[0]
    out_d = torch.zeros((n, 1056), dtype=torch.uint8, device=dev)
    res_d = torch.zeros(n, dtype=torch.int32, device=dev)
    for crc, title in ((pkg.CRC_NONE, "decode only, fixed 6 iterations"), (pkg.CRC24B, "decode only, CRC24B early stop")):
        descs = pkg.make_descs(n, 1, 384, crc_poly=crc)
        dec.decode_batch_device(descs, llr_d.data_ptr(), out_d.data_ptr(), res_d.data_ptr())
        dec.synchronize()
        assert L.ldpc_cuda_debug_phase_cycles(out) == 0
        dec.decode_batch_device(descs, llr_d.data_ptr(), out_d.data_ptr(), res_d.data_ptr())
        dec.synchronize()
        report(title)


if __name__ == "__main__":
    main()
