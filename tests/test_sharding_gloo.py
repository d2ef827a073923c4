"""The N > 1 host path on CPU: two gloo ranks partition one batch by join group, each runs the (oracle) pipeline on
its shard, the counter blocks are sum-allreduced, and the result equals the single-rank run.  The GPU kernels are
not involved here (no GPU in this container); what is under test is the partition rule, the counter layout and
the reduction — the same code bench.py runs over NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200 import sharding, synth


def _oracle_counters(w):
    from oracle import cpu as O
    j = O.process_batch(w["spans"], w["events"], w["idents"], indexed=True, nthreads=2)
    ni = len(w["inc_actual"])
    sig = O.incident_features(j["attr"], w["spans"]["incident"], ni)
    post, top1, _ = O.attribute(O.default_model(), sig)
    c = A.Counters()
    c.debug = j["debug"]
    c.confusion = O.confusion(w["inc_actual"], top1, w["inc_expected"], w["inc_expected_n"], post, 0.10, 0.01)
    c.events, c.spans, c.incidents = len(w["events"]), len(w["spans"]), ni
    c.candidates, c.kept = int(j["n_cand"].sum()), int((j["cand_event"] != A.NONE).sum())
    return c, j


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # weak-scaling layout of bench.py: every rank generates its own shard of distinct pods
    w = synth.generate_host(synth.config(20_000, 2_000, 16, 8, 2, seed=42, horizon_ns=4_000_000_000, rank=rank, world=world))
    c, _ = _oracle_counters(w)
    t = torch.from_numpy(sharding.counters_to_array(c))
    sharding.allreduce_counters(t)
    if rank == 0:
        np.save(out, t.numpy())
    dist.destroy_process_group()


def test_two_rank_counter_allreduce(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "reduced.npy")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = sharding.array_to_counters(np.load(out))
    # the same two shards on one rank
    want = np.zeros(A.COUNTER_WORDS, np.int64)
    for r in range(2):
        w = synth.generate_host(synth.config(20_000, 2_000, 16, 8, 2, seed=42, horizon_ns=4_000_000_000, rank=r, world=2))
        want += sharding.counters_to_array(_oracle_counters(w)[0])
    assert np.array_equal(sharding.counters_to_array(got), want)
    assert got.events == 40_000 and got.spans == 4_000 and got.confusion.n == got.incidents == 400
    r = sharding.ratios(got)
    assert 0.0 < r["accuracy"] <= r["partial"] <= 1.0 and 0.0 < r["coverage"] <= 1.0


def test_partition_by_group_is_exact():
    """Splitting one batch by join group and summing the shards' counters gives the unsharded result, and every
    span keeps exactly its candidates (mapped back through the shard's event index)."""
    w = synth.generate_host(synth.config(60_000, 6_000, 32, 8, 2, seed=7, horizon_ns=6_000_000_000))
    whole, j = _oracle_counters(w)
    parts = sharding.partition(w["spans"], w["events"], w["idents"], 3)
    assert sum(len(p["spans"]) for p in parts) == len(w["spans"]) and min(len(p["spans"]) for p in parts) > 0
    from oracle import cpu as O
    low = kept = 0
    for p in parts:
        pj = O.process_batch(p["spans"], p["events"], w["idents"], indexed=True)
        low += pj["debug"].low_confidence
        ce = np.where(pj["cand_event"] == A.NONE, A.NONE, p["event_index"][np.minimum(pj["cand_event"], max(len(p["event_index"]) - 1, 0))])
        assert np.array_equal(ce.astype(np.uint32), j["cand_event"][p["span_index"]])
        kept += int((pj["cand_event"] != A.NONE).sum())
    assert low == whole.debug.low_confidence and kept == whole.kept
    own = sharding.shard_of_group(np.arange(1, 4097), 8)
    assert np.bincount(own, minlength=8).min() > 400           # even spread of consecutive group ids


def _exchange_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # one batch, arriving unsharded: rank r holds every world-th record
    w = synth.generate_host(synth.config(40_000, 4_000, 32, 8, 2, seed=9, horizon_ns=5_000_000_000))
    owner = sharding.shard_of_group(w["idents"]["grp"], world)
    parts = {}
    for name, rec in (("spans", w["spans"]), ("events", w["events"])):
        mine = np.ascontiguousarray(rec[rank::world])
        t = torch.from_numpy(mine.view(np.int64).reshape(len(mine), 4).copy())
        got, src, idx = sharding.exchange(t, torch.from_numpy(owner[mine["ident"]]))
        parts[name] = got.numpy().reshape(-1).view(rec.dtype)
        parts[name + "_gidx"] = (idx.numpy() * world + src.numpy())     # position in the unsharded batch
        assert (owner[parts[name]["ident"]] == rank).all()              # every record landed on its owner
    from oracle import cpu as O
    j = O.process_batch(parts["spans"], parts["events"], w["idents"], indexed=True)
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), span_gidx=parts["spans_gidx"], event_gidx=parts["events_gidx"],
             cand=j["cand_event"], low=np.array([j["debug"].low_confidence]))
    dist.destroy_process_group()


def test_all_to_all_exchange_of_unsharded_input(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_exchange_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    w = synth.generate_host(synth.config(40_000, 4_000, 32, 8, 2, seed=9, horizon_ns=5_000_000_000))
    from oracle import cpu as O
    whole = O.process_batch(w["spans"], w["events"], w["idents"], indexed=True)
    low, seen = 0, 0
    for r in range(2):
        z = np.load(os.path.join(str(tmp_path), "rank%d.npz" % r))
        cand = np.where(z["cand"] == A.NONE, A.NONE, z["event_gidx"][np.minimum(z["cand"], len(z["event_gidx"]) - 1)]).astype(np.uint32)
        assert np.array_equal(cand, whole["cand_event"][z["span_gidx"]])
        low += int(z["low"][0])
        seen += len(z["span_gidx"])
    assert seen == 4_000 and low == whole["debug"].low_confidence
