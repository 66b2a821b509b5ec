"""How fast can pinned host memory reach the device on this box: one stream, two streams, chunk sizes (development aid)."""
import time
import torch

n = 1664 * 1024 * 1024
h = torch.empty(n, dtype=torch.int8).pin_memory()
d = torch.empty(n, dtype=torch.int8, device="cuda")
torch.cuda.synchronize()


def run(nstreams, chunk):
    streams = [torch.cuda.Stream() for _ in range(nstreams)]
    best = 1e9
    for _ in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        k = 0
        for o in range(0, n, chunk):
            with torch.cuda.stream(streams[k % nstreams]):
                d[o:o + chunk].copy_(h[o:o + chunk], non_blocking=True)
            k += 1
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return n / best / 1e9


for ns in (1, 2, 3):
    for chunk in (8 << 20, 52 << 20, 256 << 20, n):
        print("streams %d chunk %4d MB: %.1f GB/s" % (ns, chunk >> 20, run(ns, chunk)), flush=True)
