import sys, time, numpy as np, torch
sys.path.insert(0, '/root/repo')
import __graft_entry__ as ge
pkg = ge.load_package()
dec = pkg.LdpcCuda(device=0, max_codeblocks=20000)
tbs, Qm, layers, nsym = 1277992, 8, 4, 42588 * 4
descs1, _ = pkg.segment(tbs, 1, Qm, layers, nsym)
C = len(descs1); n_tb = 128
E = descs1["rm_length"].astype(np.int64); tb_bytes = int(E.sum()); off = np.concatenate([[0], np.cumsum(E)[:-1]])
d = np.zeros(n_tb * C, pkg.CB_DESC)
for i in range(n_tb):
    s = d[i*C:(i+1)*C]; s[:] = descs1
    s["crc_poly"], s["early_stop"], s["new_data"] = pkg.CRC24B, 1, 1
    s["in_offset"] = i * tb_bytes + off
    s["soft_offset"] = (i * C + np.arange(C)) * 66 * 384
    s["out_offset"] = (i * C + np.arange(C)) * 1056
dec.softbuffer_reserve(n_tb * C * 66 * 384)
rm = torch.zeros(n_tb * tb_bytes, dtype=torch.int8, device='cuda')
out = torch.zeros(n_tb * C * 1056, dtype=torch.uint8, device='cuda')
res = torch.zeros(n_tb * C, dtype=torch.int32, device='cuda')
for _ in range(3):
    dec.dematch_decode_batch_device(d, rm.data_ptr(), out.data_ptr(), res.data_ptr())
dec.synchronize()
ts = []
for _ in range(5):
    t0 = time.perf_counter()
    dec.dematch_decode_batch_device(d, rm.data_ptr(), out.data_ptr(), res.data_ptr())
    t1 = time.perf_counter()
    dec.synchronize()
    t2 = time.perf_counter()
    ts.append((t1 - t0, t2 - t0))
print("enqueue (CPU prep) ms, total ms:", [(round(a*1e3,3), round(b*1e3,3)) for a, b in ts])
