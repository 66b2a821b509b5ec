"""N > 1 host logic on CPU: two processes over gloo shard the codeblocks of transport blocks with the product's sharding
helper (srsran_project_23.5_b200/sharding.py), "decode" their share, and rank 0 reassembles the transport block.

There is no GPU here and the product has no CPU path, so the per-codeblock decode inside each worker is done by the oracle
(test infrastructure standing in for the device). What is under test is everything around it: the partition is disjoint
and complete, a HARQ process maps every codeblock to the same rank in every transmission (soft-buffer affinity), local
descriptor layouts are dense, and the gather puts every codeblock back at its place with no data-path collective."""
import os
import socket
import sys

import numpy as np
import pytest

import __graft_entry__ as ge

pkg = ge.load_package()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _sharding():
    import importlib
    return importlib.import_module(pkg.__name__ + ".sharding")


def test_partition_is_disjoint_complete_and_stable():
    S = _sharding()
    descs, _ = pkg.segment(1277992, 1, 8, 4, 42588 * 4)
    for world in (1, 2, 4, 8):
        for policy in ("round_robin", "by_process"):
            for proc in [(0, 0x4601, 3), (5, 17, 0), (15, 65535, 15)]:
                parts = [S.shard_descs(descs, r, world, *proc, policy=policy) for r in range(world)]
                allidx = np.sort(np.concatenate(parts))
                assert (allidx == np.arange(len(descs))).all()
                again = [S.shard_descs(descs, r, world, *proc, policy=policy) for r in range(world)]
                assert all((a == b).all() for a, b in zip(parts, again))
                if policy == "round_robin" and world > 1:
                    sizes = [len(p) for p in parts]
                    assert max(sizes) - min(sizes) <= 1
                if policy == "by_process":
                    assert sum(len(p) > 0 for p in parts) == 1
    # processes spread over the ranks
    homes = {S.process_home(c, r, h, 8) for c in range(4) for r in range(40) for h in range(16)}
    assert homes == set(range(8))


def test_local_layout_is_dense():
    S = _sharding()
    descs, _ = pkg.segment(200000, 1, 6, 2, 30000)
    off = 0
    for d in descs:
        d["in_offset"] = off
        off += int(d["rm_length"])
    mine = S.shard_descs(descs, 1, 2, 0, 1, 2)
    loc, src, out_off = S.local_layout(descs, mine, fused=True)
    assert len(loc) == len(mine) and int(loc["in_offset"][0]) == 0
    for k in range(1, len(loc)):
        assert int(loc["in_offset"][k]) == int(loc["in_offset"][k - 1]) + int(loc["rm_length"][k - 1])
        assert out_off[k] > out_off[k - 1]
    assert [s[0] for s in src] == [int(descs["in_offset"][i]) for i in mine]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    import oracles as O
    S = _sharding()
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        bg, Qm, layers, nsym, tbs = 2, 2, 1, 30000, 3000 * 8 - 24
        rng = np.random.default_rng(77)  # same stream in both ranks: same transport block and channel
        descs, crc_bits = pkg.segment(tbs, bg, Qm, layers, nsym)
        C = len(descs)
        proc = (3, 0x4601, 5)
        mine = S.shard_descs(descs, rank, world, *proc)
        softs = {int(i): np.zeros(50 * int(descs["lifting_size"][i]), np.int8) for i in mine}
        msgs, cws = [], []
        for r in range(C):
            Z, F = int(descs["lifting_size"][r]), int(descs["nof_filler_bits"][r])
            msg, cw = O.make_codeword(rng, bg, Z, F, O.CRC24B if C > 1 else O.CRC16)
            msgs.append(msg)
            cws.append(cw)
        done = None
        for t, rv in enumerate([0, 2]):
            again = S.shard_descs(descs, rank, world, *proc)
            assert (again == mine).all()  # HARQ affinity: the soft buffers of `mine` live here
            bits, res = [], np.zeros(len(mine), pkg.CB_RESULT)
            rx = []
            for r in range(C):  # every rank draws the whole channel so that the random streams stay aligned
                Z, F, E = int(descs["lifting_size"][r]), int(descs["nof_filler_bits"][r]), int(descs["rm_length"][r])
                rx.append(O.awgn_llrs(rng, O.port().rate_match(cws[r], E, rv, Qm, 0, F), -3.0 if t == 0 else 0.0))
            for k, i in enumerate(mine):
                i = int(i)
                Z, F = int(descs["lifting_size"][i]), int(descs["nof_filler_bits"][i])
                O.port().rate_dematch(softs[i], rx[i], t == 0, rv, Qm, 0, F)
                b, it = O.port().decode(bg, Z, softs[i], fillers=F, crc=O.CRC24B if C > 1 else O.CRC16)
                bits.append(b)
                res[k]["crc_ok"], res[k]["iterations"] = int(it > 0), it if it else 6
            all_bits, all_res = S.gather_transport_block(dist, bits, res, mine, C, dst=0)
            if rank == 0:
                assert len(all_bits) == C
                ok = bool(all_res["crc_ok"].all())
                if ok:
                    for r in range(C):
                        assert (all_bits[r] == O.msg_to_packed(msgs[r])).all(), r
                done = (t, ok, [int(x) for x in all_res["iterations"]])
        q.put((rank, len(mine), C, done))
    finally:
        dist.destroy_process_group()


def test_two_ranks_over_gloo_decode_one_transport_block():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    out.sort()
    (r0, n0, C, done), (r1, n1, _, _) = out
    assert C >= 4 and n0 + n1 == C and abs(n0 - n1) <= 1
    assert done is not None and done[1], "the transport block must decode after the retransmission"
