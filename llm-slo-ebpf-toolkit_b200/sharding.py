"""Multi-GPU host logic: records are partitioned by node/pod identity, every rank runs the whole
single-GPU pipeline on its shard, and the only exchange is a sum-allreduce of the integer counter block
(lsg_counters).  Ratios are formed from the reduced integers, so they do not depend on the reduction order.

Nothing here computes the path; it only decides which rank owns which record and reduces counters through
torch.distributed (NCCL on GPUs, gloo in the CPU tests)."""
import ctypes as C

import numpy as np

from . import _abi as A


def shard_of_group(grp, world):
    """Owner rank of a join group.  Groups are closed under every relation Match() can use (same pod, same
    (service, node), same trace id), so sending a whole group to one rank keeps the join shard-local."""
    g = np.asarray(grp, dtype=np.uint64)
    h = (g * np.uint64(0x9E3779B97F4A7C15)) >> np.uint64(33)     # Fibonacci hashing: consecutive group ids spread evenly
    out = (h % np.uint64(world)).astype(np.int64)
    out[np.asarray(grp) == 0] = 0                                 # group 0 matches nothing: park it on rank 0
    return out


def partition(spans, events, idents, world):
    """Split one interned batch into per-rank (spans, events, span_index, event_index).  The identity table is
    replicated (it is tiny); indices map shard-local results back to the caller's order."""
    owner = shard_of_group(idents["grp"], world)
    s_owner, e_owner = owner[spans["ident"]], owner[events["ident"]]
    parts = []
    for r in range(world):
        si, ei = np.nonzero(s_owner == r)[0], np.nonzero(e_owner == r)[0]
        parts.append({"spans": spans[si], "events": events[ei], "span_index": si, "event_index": ei})
    return parts


def counters_to_array(c):
    arr = np.zeros(A.COUNTER_WORDS, np.int64)
    C.memmove(arr.ctypes.data, C.byref(c), C.sizeof(c))
    return arr


def array_to_counters(arr):
    c = A.Counters()
    arr = np.ascontiguousarray(arr, dtype=np.int64)
    C.memmove(C.byref(c), arr.ctypes.data, C.sizeof(c))
    return c


def allreduce_counters(counters_tensor):
    """In-place sum over the default process group; a no-op without one."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(counters_tensor, op=dist.ReduceOp.SUM)
    return counters_tensor


def ratios(c):
    """Accuracy, partial accuracy and coverage accuracy (pipeline.go:85-185) from reduced integers."""
    n = int(c.confusion.n)
    if n == 0:
        return {"accuracy": 0.0, "partial": 0.0, "coverage": 0.0}
    cov = 0.0
    for f in range(A.NDOM + 1):
        for ln in range(1, A.NDOM + 1):
            cov += int(c.confusion.coverage[f][ln]) * (float(f) / float(ln))
    return {"accuracy": c.confusion.correct / n, "partial": c.confusion.partial / n, "coverage": cov / n}


def exchange(records, owner, group=None):
    """All-to-all of packed records for input that arrives unsharded (the optional pre-step): every
    rank holds an arbitrary slice of the batch; `owner[i]` is the rank that must process records[i] (shard_of_group of
    its join group).  Returns (received records, source rank of each, index of each in its source slice).
    `records` is a torch tensor [n, k] of any dtype (32-byte records as 4 x int64 or 32 x uint8) on the device the
    process group works on (CUDA for NCCL, CPU for gloo)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    owner = owner.to(torch.int64)
    order = torch.argsort(owner, stable=True)
    send = records[order].contiguous()
    send_idx = order.to(torch.int64)
    counts = torch.bincount(owner, minlength=world).to(torch.int64)
    recv_counts = torch.empty_like(counts)
    dist.all_to_all_single(recv_counts, counts, group=group)
    sc, rc = counts.tolist(), recv_counts.tolist()
    out = records.new_empty((int(sum(rc)),) + tuple(records.shape[1:]))
    dist.all_to_all_single(out, send, output_split_sizes=rc, input_split_sizes=sc, group=group)
    idx = send_idx.new_empty(int(sum(rc)))
    dist.all_to_all_single(idx, send_idx, output_split_sizes=rc, input_split_sizes=sc, group=group)
    src = torch.repeat_interleave(torch.arange(world, device=idx.device), torch.tensor(rc, device=idx.device))
    return out, src, idx
