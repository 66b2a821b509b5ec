"""Multi-GPU sharding of the LDPC receive chain: which rank decodes which codeblock.

Codeblocks are independent (the loop body of pusch_decoder_impl::decode, pusch_decoder_impl.cpp:156-201, shares nothing
between codeblocks except the transport-block offsets), so the path shards with NO data-path collective:

* new-data codeblocks of a slot go round-robin, codeblock i -> rank i mod world;
* every codeblock of a (cell, rnti, harq_id) process goes where that process' soft buffers live: the rank that decoded
  its first transmission. Two policies give that: `round_robin` remembers nothing and spreads the codeblocks of a TB over
  all ranks (the soft buffer of codeblock r of a process always lives on rank (first + r) mod world, `first` a stable
  hash of the process, so a retransmission finds it without any table), `by_process` keeps a whole TB on one rank.

The only rank-to-rank traffic is the control-plane gather of per-codeblock results ({crc_ok, iterations}, 4 bytes each) and
of the decoded bytes to the rank that assembles the transport block; `gather_transport_block` does it with
torch.distributed (NCCL on GPUs, gloo in the CPU tests). Nothing here touches LLRs or soft buffers.
"""
import zlib

import numpy as np


def process_home(cell, rnti, harq_id, world):
    """Stable rank of a HARQ process (same value in every rank and every run; Python's hash() is salted)."""
    key = np.array([cell, rnti, harq_id], np.uint32).tobytes()
    return zlib.crc32(key) % world


def codeblock_rank(cb_index, world, cell=0, rnti=0, harq_id=0, policy="round_robin"):
    """Rank that owns codeblock cb_index of the transport block of HARQ process (cell, rnti, harq_id)."""
    home = process_home(cell, rnti, harq_id, world)
    if policy == "by_process":
        return home
    if policy == "round_robin":
        return (home + cb_index) % world
    raise ValueError("unknown sharding policy %r" % policy)


def shard_descs(descs, rank, world, cell=0, rnti=0, harq_id=0, policy="round_robin"):
    """Indices (into descs) of the codeblocks this rank decodes, in transport-block order."""
    n = len(descs)
    if world == 1:
        return np.arange(n)
    owner = np.array([codeblock_rank(i, world, cell, rnti, harq_id, policy) for i in range(n)])
    return np.nonzero(owner == rank)[0]


def local_layout(descs, mine, fused):
    """Descriptors of this rank's codeblocks with dense local offsets (input, output and soft buffer) and the byte ranges
    of the global input they came from. Returns (local descs, [(global in offset, length)], local out offsets)."""
    loc = descs[mine].copy()
    in_off = out_off = 0
    src = []
    for k in range(len(loc)):
        d = loc[k]
        length = int(d["rm_length"] if fused else d["in_length"])
        src.append((int(d["in_offset"]), length))
        kb = ((10 if int(d["base_graph"]) == 2 else 22) * int(d["lifting_size"]) + 7) // 8
        loc[k]["in_offset"], loc[k]["out_offset"] = in_off, out_off
        in_off += length
        out_off += kb
    return loc, src, [int(x) for x in loc["out_offset"]]


def gather_transport_block(dist, local_bits, local_results, mine, n_total, dst=0):
    """Collects per-codeblock results and decoded bytes on rank dst (control plane; a few KB per TB).

    local_bits: list of uint8 arrays (one per local codeblock), local_results: structured array of ldpc_cuda_cb_result.
    Returns (list of n_total byte arrays, results[n_total]) on dst, (None, None) elsewhere."""
    rank, world = dist.get_rank(), dist.get_world_size()
    payload = (np.asarray(mine), [np.asarray(b) for b in local_bits], np.asarray(local_results))
    gathered = [None] * world if rank == dst else None
    dist.gather_object(payload, gathered, dst=dst)
    if rank != dst:
        return None, None
    bits = [None] * n_total
    results = np.zeros(n_total, payload[2].dtype)
    for idx, bl, res in gathered:
        for k, i in enumerate(idx):
            bits[int(i)] = bl[k]
            results[int(i)] = res[k]
    assert all(b is not None for b in bits), "a codeblock was decoded by no rank"
    return bits, results
