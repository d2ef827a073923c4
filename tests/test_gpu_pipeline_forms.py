"""lsg_pipeline with the two optional forms added in round 2, through the C ABI with host buffers:

events24      the 24-byte upload form of the events (include/llmslo_gpu.h: lsg_event24).  Every output and every
              counter must equal the run on the 32-byte records, on both implementations of the sort + join, through
              the blocking call and through submit / wait; a base that does not cover the spans' range is refused.
features_out  count / sum / log2 sketch per (incident, signal) accumulated in the join's epilogue.  No reference
              counterpart (only the maxima have one, correlator.go:109-111: parity unpinned); checked against the
              oracle's restatement on the same enriched attributes: counts and sketch bins exactly, sums to 1e-9."""
import ctypes as C

import numpy as np
import pytest

import llm_slo_ebpf_toolkit_b200 as pkg
from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200 import synth
from oracle import cpu as O

pytestmark = pytest.mark.gpu
S = 1_000_000_000


def _args(w, fanout=3, events24=None, base=0, features=False):
    n_s, n_i = len(w["spans"]), len(w["inc_actual"])
    keep = {k: np.ascontiguousarray(w[k]) for k in ("spans", "events", "idents", "inc_actual", "inc_expected", "inc_expected_n")}
    o = {"ce": np.zeros(n_s * fanout, np.uint32), "ct": np.zeros(n_s * fanout, np.uint8), "attr": np.zeros(n_s * A.NSIG),
         "conf": np.zeros(n_s), "sig": np.zeros(n_i * A.NSIG), "post": np.zeros(n_i * A.NDOM), "top1": np.zeros(n_i, np.uint8),
         "feat": np.zeros(n_i * A.NSIG, A.FEATURE_DTYPE)}
    args = A.PipelineArgs(keep["spans"].ctypes.data, n_s, None if events24 is not None else keep["events"].ctypes.data,
                          len(w["events"]), keep["idents"].ctypes.data, len(w["idents"]), n_i, keep["inc_actual"].ctypes.data,
                          keep["inc_expected"].ctypes.data, keep["inc_expected_n"].ctypes.data, 0, 0.0, fanout, 0.10,
                          o["ce"].ctypes.data, o["ct"].ctypes.data, o["attr"].ctypes.data, o["conf"].ctypes.data,
                          o["sig"].ctypes.data, o["post"].ctypes.data, o["top1"].ctypes.data, None, None, None)
    if events24 is not None:
        args.events24, args.events24_base = events24.ctypes.data, int(base)
    if features:
        args.features_out = o["feat"].ctypes.data
    return args, (keep, events24), o


def _same(a, b):
    for k in ("ce", "ct", "attr", "conf", "sig", "post", "top1"):
        assert np.array_equal(a[k], b[k], equal_nan=True), k


@pytest.mark.parametrize("path", [0, 1])
def test_events24_equals_the_32_byte_form(ctx, path):
    lib = pkg.load_library()
    cfg = synth.config(300_000, 30_000, 64, 8, 2, seed=21, horizon_ns=120 * S)
    w = synth.generate_host(cfg)
    ev = w["events"]
    rng = np.random.default_rng(2)
    ev["ts_ns"][rng.choice(len(ev), 500, replace=False)] = A.TS_ZERO
    ev["ts_ns"][rng.choice(len(ev), 500, replace=False)] += 90 * 86400 * S      # far outside: matches nothing either way
    ev["trace_sig"][rng.choice(len(ev), 300, replace=False)] |= 15               # unsupported types are still counted
    base = int(w["spans"]["ts_ns"].min()) - 2 * S
    packed = np.zeros(len(ev), A.EVENT24_DTYPE)
    assert lib.lsg_pack_events24(ev.ctypes.data, len(ev), base, packed.ctypes.data) == 0
    ctx.set_join_path(path)
    try:
        a32, k32, o32 = _args(w)
        c32 = A.Counters()
        ctx.call_dev("lsg_pipeline", C.byref(a32), C.byref(c32))
        a24, k24, o24 = _args(w, events24=packed, base=base)
        c24 = A.Counters()
        ctx.call_dev("lsg_pipeline", C.byref(a24), C.byref(c24))
        assert bytes(c24) == bytes(c32) and c24.kept > 10_000
        _same(o24, o32)
        # double-buffered form
        b24, kb, ob = _args(w, events24=packed, base=base)
        ctx.call_dev("lsg_pipeline_submit", C.byref(b24), 1)
        cb = A.Counters()
        ctx.call_dev("lsg_pipeline_wait", 1, C.byref(cb))
        assert bytes(cb) == bytes(c32)
        _same(ob, o32)
        # a base after the first span - window: the offsets cannot express every time that may match -> refused
        bad, kk, oo = _args(w, events24=packed, base=base + 3 * S)
        with pytest.raises(pkg.LsgError) as e:
            ctx.call_dev("lsg_pipeline", C.byref(bad), C.byref(A.Counters()))
        assert e.value.code == -4
    finally:
        ctx.set_join_path(-1)


def test_widen_kernel_equals_host_unpack(ctx):
    import torch
    lib = pkg.load_library()
    cfg = synth.config(100_003, 10_000, 64, 8, 2, seed=5)
    ev = synth.generate_host(cfg)["events"]
    base = int(cfg.t0_ns) - 2 * S
    packed = np.zeros(len(ev), A.EVENT24_DTYPE)
    assert lib.lsg_pack_events24(ev.ctypes.data, len(ev), base, packed.ctypes.data) == 0
    want = np.zeros(len(ev), A.EVENT_DTYPE)
    lib.lsg_unpack_events24(packed.ctypes.data, len(ev), base, want.ctypes.data)
    d_in = torch.from_numpy(packed.view(np.uint8)).cuda()
    d_out = torch.empty(len(ev) * 32, dtype=torch.uint8, device="cuda")
    ctx.call_dev("lsg_widen_events24_dev", d_in.data_ptr(), len(ev), base, d_out.data_ptr())
    assert d_out.cpu().numpy().tobytes() == want.tobytes()
    inside = ev["ts_ns"] >= base
    assert want[inside].tobytes() == ev[inside].tobytes()


@pytest.mark.parametrize("path", [0, 1])
def test_pipeline_features_out(ctx, path):
    cfg = synth.config(200_000, 20_000, 16, 8, 2, seed=33, horizon_ns=40 * S, multi_fault=1)
    w = synth.generate_host(cfg)
    ctx.set_join_path(path)
    try:
        args, keep, o = _args(w, features=True)
        c = A.Counters()
        ctx.call_dev("lsg_pipeline", C.byref(args), C.byref(c))
    finally:
        ctx.set_join_path(-1)
    n_i = len(w["inc_actual"])
    want_sig, want_feat = O.incident_features(o["attr"].reshape(-1, A.NSIG), w["spans"]["incident"], n_i, want_features=True)
    assert np.array_equal(o["sig"].reshape(-1, A.NSIG), want_sig, equal_nan=True)
    got = o["feat"].reshape(n_i, A.NSIG)
    assert np.array_equal(got["count"], want_feat["count"]) and got["count"].sum() > 10_000
    assert np.array_equal(got["sketch"], want_feat["sketch"])
    assert np.allclose(got["sum"], want_feat["sum"], rtol=1e-9, atol=0)
