"""lsg_multi_*: the path on every GPU of the box behind the C ABI.  One context per device, one
shard of distinct pods per rank, ncclAllReduce(sum) of the counter block inside the library.

The reduced counters must equal what ONE context returns for the concatenated batch (the reference's ProcessBatch +
BuildConfusionMatrix on everything at once: processor.go:29-58, pipeline.go:65-82), and every rank's per-span /
per-incident outputs must equal the single-context run on its shard.  Runs with as many ranks as the box has GPUs
(1 on a one-GPU box: the communicator then has a single rank, the code path is the same)."""
import ctypes as C

import numpy as np
import pytest

import llm_slo_ebpf_toolkit_b200 as pkg
from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200 import synth

pytestmark = pytest.mark.gpu
S = 1_000_000_000


def _host_args(w, fanout=3):
    n_s, n_i = len(w["spans"]), len(w["inc_actual"])
    keep = {k: np.ascontiguousarray(w[k]) for k in ("spans", "events", "idents", "inc_actual", "inc_expected", "inc_expected_n")}
    o = {"ce": np.zeros(n_s * fanout, np.uint32), "ct": np.zeros(n_s * fanout, np.uint8), "attr": np.zeros(n_s * A.NSIG),
         "conf": np.zeros(n_s), "sig": np.zeros(n_i * A.NSIG), "post": np.zeros(n_i * A.NDOM), "top1": np.zeros(n_i, np.uint8)}
    args = A.PipelineArgs(keep["spans"].ctypes.data, n_s, keep["events"].ctypes.data, len(w["events"]), keep["idents"].ctypes.data,
                          len(w["idents"]), n_i, keep["inc_actual"].ctypes.data, keep["inc_expected"].ctypes.data,
                          keep["inc_expected_n"].ctypes.data, 0, 0.0, fanout, 0.10, o["ce"].ctypes.data, o["ct"].ctypes.data,
                          o["attr"].ctypes.data, o["conf"].ctypes.data, o["sig"].ctypes.data, o["post"].ctypes.data,
                          o["top1"].ctypes.data, None, None, None)
    return args, keep, o


def _concatenate(shards):
    """One batch out of the shards: identity rows, groups, trace ids and incidents renumbered so that nothing collides."""
    out = {k: [] for k in shards[0]}
    id_off = grp_off = tr_off = inc_off = 0
    for w in shards:
        ev, sp, ids = w["events"].copy(), w["spans"].copy(), w["idents"].copy()
        ev["ident"] += id_off
        sp["ident"] += id_off
        tr = ev["trace_sig"] >> 4
        ev["trace_sig"] = (np.where(tr != 0, tr + tr_off, 0).astype(np.uint32) << 4) | (ev["trace_sig"] & 15)
        sp["trace"] = np.where(sp["trace"] != 0, sp["trace"] + tr_off, 0)
        sp["incident"] = np.where(sp["incident"] != A.NONE, sp["incident"] + inc_off, A.NONE)
        ids["grp"] = np.where(ids["grp"] != 0, ids["grp"] + grp_off, 0)
        for k, v in (("events", ev), ("spans", sp), ("idents", ids), ("inc_actual", w["inc_actual"]),
                     ("inc_expected", w["inc_expected"]), ("inc_expected_n", w["inc_expected_n"])):
            out[k].append(v)
        id_off += len(ids)
        grp_off += int(ids["grp"].max())
        tr_off += int(w["spans"]["trace"].max())
        inc_off += len(w["inc_actual"])
    return {k: np.concatenate(v) for k, v in out.items()}


def test_multi_pipeline_equals_one_context_on_the_whole_batch(ctx):
    import torch
    lib = pkg.load_library()
    n = torch.cuda.device_count()
    shards = [synth.generate_host(synth.config(150_000 + 10_000 * r, 15_000 + 1_000 * r, 16, 8, 2, seed=50, horizon_ns=40 * S,
                                               multi_fault=1, rank=r, world=n)) for r in range(n)]
    for r, w in enumerate(shards):     # unsupported types and dead records on every shard, in different numbers
        w["events"]["trace_sig"][: 100 * (r + 1)] |= 15
        w["events"]["ts_ns"][200: 200 + 50 * (r + 1)] = A.TS_ZERO
    h = C.c_void_p()
    devs = (C.c_int * n)(*range(n))
    rc = lib.lsg_multi_create(devs, n, C.byref(h))
    assert rc == 0, lib.lsg_multi_last_error(h)
    try:
        assert lib.lsg_multi_size(h) == n and lib.lsg_multi_ctx(h, 0) and not lib.lsg_multi_ctx(h, n)
        per = [_host_args(w) for w in shards]
        arr = (A.PipelineArgs * n)(*[p[0] for p in per])
        own = (A.Counters * n)()
        red = A.Counters()
        for _ in range(2):             # twice: the handle is reusable
            rc = lib.lsg_multi_pipeline(h, arr, own, C.byref(red))
            assert rc == 0, lib.lsg_multi_last_error(h)
        # every rank against the single-context run on its shard
        for r, w in enumerate(shards):
            a1, k1, o1 = _host_args(w)
            c1 = A.Counters()
            ctx.call_dev("lsg_pipeline", C.byref(a1), C.byref(c1))
            assert bytes(own[r]) == bytes(c1)
            for k in o1:
                assert np.array_equal(per[r][2][k], o1[k], equal_nan=True), (r, k)
        # the reduced block against one context on the concatenated batch
        whole = _concatenate(shards)
        aw, kw, ow = _host_args(whole)
        cw = A.Counters()
        ctx.call_dev("lsg_pipeline", C.byref(aw), C.byref(cw))
        assert bytes(red.confusion) == bytes(cw.confusion)
        assert red.debug.as_tuple() == cw.debug.as_tuple()
        assert (red.events, red.spans, red.incidents, red.candidates, red.kept, red.unsupported_events, red.bad_idents) == \
            (cw.events, cw.spans, cw.incidents, cw.candidates, cw.kept, cw.unsupported_events, 0)
        assert red.kept > 10_000 * n
        assert red.unsupported_events == sum(int(((w["events"]["trace_sig"] & 15) >= A.NSIG).sum()) for w in shards) > 100 * n
        # a bad shard is reported with its rank, and the handle stays usable
        bad = shards[0]["events"].copy()
        bad["ident"][7] = 10_000
        wb = dict(shards[0], events=bad)
        pb = _host_args(wb)
        arr_bad = (A.PipelineArgs * n)(*([pb[0]] + [p[0] for p in per[1:]]))
        assert lib.lsg_multi_pipeline(h, arr_bad, None, C.byref(red)) == -4
        assert b"rank 0" in lib.lsg_multi_last_error(h)
        assert lib.lsg_multi_pipeline(h, arr, own, C.byref(red)) == 0
    finally:
        lib.lsg_multi_destroy(h)


def test_multi_create_rejects_bad_device_lists():
    lib = pkg.load_library()
    h = C.c_void_p()
    assert lib.lsg_multi_create((C.c_int * 2)(0, 0), 2, C.byref(h)) == -1 and not h.value      # one rank per device
    assert lib.lsg_multi_create((C.c_int * 1)(0), 0, C.byref(h)) == -1
    rc = lib.lsg_multi_create((C.c_int * 1)(99), 1, C.byref(h))
    assert rc == -5 and h.value
    assert b"device 99" in lib.lsg_multi_last_error(h)
    lib.lsg_multi_destroy(h)
