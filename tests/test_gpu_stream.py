"""GPU parity for the rows either side of the path: F1 retry-storm counts, F2 ring-buffer decode,
F4 signal status.  Vectors transcribed from pkg/correlation/retry_storm_test.go and pkg/collector/ringbuf_test.go."""
import struct

import numpy as np
import pytest

from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200 import correlation as CO
from llm_slo_ebpf_toolkit_b200 import ringbuf as RB
from llm_slo_ebpf_toolkit_b200 import synth
from oracle import cpu as O
from oracle import ref_str as R

pytestmark = pytest.mark.gpu
NOW = 1_771_372_800_000_000_000
S = CO.SECOND


def test_retry_storm_reference_tests(ctx):  # retry_storm_test.go:8-110
    d = CO.NewRetryStormDetector(10 * S, 5, ctx)
    for i in range(4):
        assert not d.Record("pod-a", NOW + i * S)
    assert not d.IsStorm("pod-a", NOW + 4 * S)
    assert d.Record("pod-a", NOW + 4 * S) and d.IsStorm("pod-a", NOW + 4 * S)
    d = CO.NewRetryStormDetector(10 * S, 5, ctx)
    assert d.RecordBatch(["pod-a"] * 5, [NOW] * 5) == [False, False, False, False, True]
    assert d.IsStorm("pod-a", NOW) and not d.IsStorm("pod-a", NOW + 11 * S) and d.Count("pod-a", NOW + 11 * S) == 0
    d = CO.NewRetryStormDetector(10 * S, 3, ctx)
    d.RecordBatch(["pod-a", "pod-a", "pod-a", "pod-b"], [NOW, NOW + S, NOW + 2 * S, NOW])
    assert d.IsStorm("pod-a", NOW + 2 * S) and not d.IsStorm("pod-b", NOW + 2 * S)
    d.Reset()
    assert not d.IsStorm("pod-a", NOW) and d.Count("pod-a", NOW) == 0
    assert not d.IsStorm("nonexistent", NOW) and d.Count("nonexistent", NOW) == 0


def test_retry_storm_batch_vs_oracle(ctx):
    cfg = synth.config(400_000, 2_000, 32, 8, 2, seed=21, horizon_ns=120 * S)
    w = synth.generate_host(cfg)
    ev, ids = w["events"], w["idents"]
    # arrival order = time order per pod (the precondition under which the reference is order-independent)
    order = np.argsort(ev["ts_ns"], kind="stable")
    ev = ev[order]
    ev["ts_ns"][::97] = A.TS_ZERO
    for window, thr in ((10 * S, 5), (S, 2), (0, 1), (3600 * S, 50)):
        cnt, storm = ctx.retry_storm(ev, ids, window, thr)
        is_rt = ((ev["trace_sig"] & 15) == 1) & (ev["ts_ns"] != A.TS_ZERO)
        pods = ids["pod"][ev["ident"]][is_rt]
        ocnt, ostorm = O.retry_storm(pods, ev["ts_ns"][is_rt], int(ids["pod"].max()) + 1, window, thr)
        assert np.array_equal(cnt[is_rt], ocnt) and np.array_equal(storm[is_rt], ostorm)
        assert not cnt[~is_rt].any() and not storm[~is_rt].any()
    assert ocnt.max() > 50
    # and against the string detector on a small prefix
    det = R.RetryStormDetector(10 * S, 5)
    cnt, storm = ctx.retry_storm(ev[:20000], ids, 10 * S, 5)
    for i in np.nonzero(((ev["trace_sig"][:20000] & 15) == 1) & (ev["ts_ns"][:20000] != A.TS_ZERO))[0]:
        assert det.Record(str(ids["pod"][ev["ident"][i]]), int(ev["ts_ns"][i])) == bool(storm[i])


def _record(pid, tid, ts, st, value, sport, dport, ip, errno):
    return struct.pack("<IIQIQHHIi", pid, tid, ts, st, value, sport, dport, ip, errno)   # binary.Write(LittleEndian, bpfEvent)


def test_ringbuf_decode_reference_vectors(ctx):  # ringbuf_test.go:9-112
    raw = _record(1234, 1235, 9999999999, 1, 5000000, 42424, 53, 0x0100007F, 0)
    assert len(raw) == RB.RECORD_BYTES
    b = RB.DecodeBatch(raw, ident=3, trace=77, ts_offset_ns=1000, ctx=ctx)
    e = b.events[0]
    assert e["pid"] == 1234 and e["ts_ns"] == 9999999999 + 1000 and e["ident"] == 3
    assert RB.signal_name(int(e["trace_sig"])) == "dns_latency_ms" and e["trace_sig"] >> 4 == 77
    assert e["value"] == 5.0 and e["conn"] != 0
    # TestSignalFromType / TestConvertValue
    cases = [(1, 5000000, 5.0), (2, 3, 3.0), (6, 50000, 50000.0), (7, 500000, 0.5), (8, 2000000, 2.0), (9, 5000000, 5.0),
             (3, 1, 1e-6), (4, 123456789, 123.456789), (5, 0, 0.0), (42, 1000000, 1.0)]
    raw = b"".join(_record(1, 1, 10 + i, st, v, 0, 0, 0, -110 if i == 2 else 0) for i, (st, v, _) in enumerate(cases))
    b = RB.DecodeBatch(raw, ctx=ctx)
    for i, (st, v, want) in enumerate(cases):
        name, _unit = RB.signalFromType(st)
        assert RB.signal_name(int(b.events[i]["trace_sig"])) == name
        assert b.events[i]["value"] == want == (float(v) if st in (2, 6) else float(v) / 1e6)
        assert b.events[i]["conn"] == 0           # no ports: no ConnTuple (ringbuf.go:184)
    assert b.errno[2] == -110 and b.errno[0] == 0
    with pytest.raises(ValueError):
        RB.DecodeBatch(raw[:-1], ctx=ctx)


def test_ringbuf_decode_bulk_and_status(ctx):
    rng = np.random.default_rng(4)
    n = 200_000
    st = rng.integers(0, 12, n).astype(np.uint32)
    val = rng.integers(0, 400_000_000, n).astype(np.uint64)
    val[st == 2] = rng.integers(0, 8, int((st == 2).sum()))
    ts = rng.integers(1, 1 << 50, n).astype(np.uint64)
    sport = rng.integers(0, 3, n).astype(np.uint16) * 1000
    dport = rng.integers(0, 2, n).astype(np.uint16) * 443
    ip = rng.integers(0, 4, n).astype(np.uint32)
    rec = np.zeros(n, dtype=np.dtype([("pid", "<u4"), ("tid", "<u4"), ("ts", "<u8"), ("st", "<u4"), ("val", "<u8"), ("sp", "<u2"),
                                      ("dp", "<u2"), ("ip", "<u4"), ("errno", "<i4")], align=False))
    assert rec.dtype.itemsize == 40
    rec["pid"], rec["ts"], rec["st"], rec["val"], rec["sp"], rec["dp"], rec["ip"] = rng.integers(1, 5000, n), ts, st, val, sport, dport, ip
    ev, status, _ = ctx.decode_ringbuf(rec.view(np.uint8).ravel(), ident=1, trace=0, ts_offset_ns=-5)
    sig_of = np.full(12, 15)
    sig_of[1:10] = [0, 1, 2, 3, 4, 5, 7, 8, 9]
    assert np.array_equal(ev["trace_sig"] & 15, sig_of[st])
    want_val = np.where((st == 2) | (st == 6), val.astype(np.float64), val.astype(np.float64) / 1e6)
    assert np.array_equal(ev["value"], want_val) and np.array_equal(ev["ts_ns"], ts.astype(np.int64) - 5)
    assert np.array_equal(ev["conn"] == 0, (sport == 0) & (dport == 0))
    tup = (sport.astype(np.uint64) << 48) | (dport.astype(np.uint64) << 32) | ip
    has = ev["conn"] != 0
    assert len(np.unique(ev["conn"][has])) == len(np.unique(tup[has]))      # equal tuples <-> equal ids on this data
    # F4: signalStatus, generator.go:203-242
    warn = np.array([40, 2, 10, 80, 60, 2, 40, 5, 10, 50, 1, 1], float)
    err = np.array([120, 5, 25, 180, 160, 8, 120, 20, 50, 200, 3, 3], float)
    sig = ev["trace_sig"] & 15
    ok = sig < 12
    want = np.zeros(n, np.uint8)
    want[ok] = (ev["value"][ok] >= warn[sig[ok]]).astype(np.uint8) + (ev["value"][ok] >= err[sig[ok]]).astype(np.uint8)
    assert np.array_equal(status, want) and np.array_equal(ctx.signal_status(ev), want)
    assert set(np.unique(want)) == {0, 1, 2}
