"""GPU parity, part 2: the CUDA path against the CPU oracle on seeded inputs through the C ABI —
randomised string batches (via the interner), the synthetic workload at sizes the oracle finishes in
seconds, edge cases, and size-independent properties at BASELINE.json's full C2 size.
Bit-exact: match sets and their order, tiers, counters, per-signal max, retrieval sums, elevated masks,
argmax labels, confusion counts.  Within 1e-6 relative: posteriors, feature sums."""
import ctypes as C
import math

import numpy as np
import pytest

from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200 import ebpfcorrelator as EC
from llm_slo_ebpf_toolkit_b200 import correlation as CO
from llm_slo_ebpf_toolkit_b200 import synth
from llm_slo_ebpf_toolkit_b200.interner import Interner
from oracle import cpu as O
from oracle import ref_str as R
from tests.randgen import random_batch

pytestmark = pytest.mark.gpu


def _eq_batch(got, want, fanout=None):
    for k in ("cand_event", "cand_tier", "conf", "retrieval", "n_cand"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(np.nan_to_num(got["attr"], nan=-1e300), np.nan_to_num(want["attr"], nan=-1e300)), "attr"
    assert got["debug"].as_tuple() == want["debug"].as_tuple()


@pytest.mark.parametrize("seed", range(8))
def test_random_string_batches_vs_string_oracle(ctx, seed):
    spans, signals = random_batch(100 + seed, 40, 300, n_pods=3 + seed % 4, dense=seed % 2 == 0)
    params = [(2000, 0.7, 3), (2000, 0.85, 1), (300, 0.7, 8), (50, 0.5, 3), (5000, 0.95, 2), (0, 0.0, 0), (2000, 0.6, 5),
              (120, 0.9, 4)][seed]
    want = R.Correlator(*params).ProcessBatch(spans, signals)
    mine = EC.Correlator(params[0], params[1], params[2], ctx).ProcessBatch(
        [EC.SpanRecord(s.TraceID, s.SpanID, s.Service, s.Node, s.Pod, s.PID, s.ConnTuple, s.Timestamp, s.Attributes)
         for s in spans],
        [CO.SignalRef(g.Signal, g.TraceID, g.Service, g.Node, g.Pod, g.PID, g.ConnTuple, g.Timestamp, g.Value) for g in signals])
    assert (mine.Debug.Unmatched, mine.Debug.LowConfidence, mine.Debug.FanoutDropped, mine.Debug.UnsupportedType) == \
        (want.Debug.Unmatched, want.Debug.LowConfidence, want.Debug.FanoutDropped, want.Debug.UnsupportedType)
    for g, w in zip(mine.Spans, want.Spans):
        assert g.Attributes == w.Attributes
        assert [c.Index for c in g.Candidates] == [c.Index for c in w._candidates]
        assert [c.Decision.Tier for c in g.Candidates] == [c.Decision.Tier for c in w._candidates]


@pytest.mark.parametrize("window_ms,thr,fanout,horizon_s,n_pods,ppg", [
    (2000, 0.7, 3, 36, 64, 2), (2000, 0.7, 3, 3600, 64, 2), (100, 0.85, 1, 36, 16, 1), (5000, 0.6, 8, 20, 8, 4),
    (2000, 1.0, 2, 36, 64, 2), (250, 0.7, 5, 4, 4, 2), (2000, 0.6, 12, 36, 16, 2), (1000, 0.7, 32, 10, 8, 2),
    (2000, 0.7, 3, 3600, 4096, 2),   # 2048 groups x 2^17 quanta: too many keys for the counting sort -> radix path
    # the counted (not enumerated) service+node tier: inner window below / between / above the tier limits, a time
    # quantum wider than the inner window, one-pod and four-pod classes, every cut below tier 4
    (50, 0.7, 3, 20, 16, 2), (400, 0.7, 3, 20, 16, 2), (20000, 0.7, 3, 60, 16, 2), (700, 0.95, 3, 20, 16, 1),
    (2000, 0.85, 4, 30, 16, 4), (499, 0.66, 2, 10, 8, 2), (501, 0.8, 3, 10, 8, 2),
])
def test_synthetic_vs_indexed_oracle(ctx, window_ms, thr, fanout, horizon_s, n_pods, ppg):
    cfg = synth.config(200_000, 20_000, n_pods, max(ppg, 8 if n_pods >= 8 else n_pods), ppg, seed=window_ms + fanout,
                       horizon_ns=horizon_s * 1_000_000_000, multi_fault=fanout % 2)
    w = synth.generate_host(cfg)
    got = ctx.process_batch(w["spans"], w["events"], w["idents"], window_ms * 1_000_000, thr, fanout)
    want = O.process_batch(w["spans"], w["events"], w["idents"], window_ms * 1_000_000, thr, fanout, indexed=True, nthreads=8)
    _eq_batch(got, want)
    assert (got["cand_event"] != A.NONE).sum() > 1000


def test_counted_tier4_with_empty_service_or_node(ctx):
    """Tiles whose events all carry one (service, node) class count the service+node tier instead of enumerating it;
    rows with an empty service or node have no class and must never match on that tier (dns.go:69-73)."""
    cfg = synth.config(120_000, 12_000, 16, 8, 2, seed=77, horizon_ns=30_000_000_000)
    w = synth.generate_host(cfg)
    for variant in range(4):
        ids = w["idents"].copy()
        if variant == 0:
            ids["service"][:] = 0                      # no row has a class: every tile is uniform with class 0
        elif variant == 1:
            ids["node"][ids["grp"] == ids["grp"].max()] = 0   # one group without a class, the others keep theirs
        elif variant == 2:
            ids["service"][::2] = 0                    # every group holds one row with and one without a class: mixed tiles
        else:
            ids["pod"][:] = 0                          # no pod anywhere: only trace ids and service+node can match
        got = ctx.process_batch(w["spans"], w["events"], ids)
        want = O.process_batch(w["spans"], w["events"], ids, indexed=True, nthreads=8)
        _eq_batch(got, want)


def test_key_range_comes_from_the_spans(ctx):
    """The time axis of the sort is the spans' range widened by the window: events far outside it (clock glitches,
    a replay file that holds a whole day) match nothing, do not widen the key space and still count as unmatched."""
    cfg = synth.config(150_000, 15_000, 16, 8, 2, seed=31, horizon_ns=30_000_000_000)
    w = synth.generate_host(cfg)
    ev = w["events"].copy()
    rng = np.random.default_rng(3)
    far = rng.choice(len(ev), 3000, replace=False)
    ev["ts_ns"][far[:1000]] += 400 * 86400 * 10**9          # more than a year later
    ev["ts_ns"][far[1000:2000]] -= 10 * 86400 * 10**9         # ten days earlier
    ev["ts_ns"][far[2000:2500]] = (1 << 62)
    ev["ts_ns"][far[2500:]] = -(1 << 62)
    got = ctx.process_batch(w["spans"], ev, w["idents"])
    want = O.process_batch(w["spans"], ev, w["idents"], indexed=True, nthreads=8)
    _eq_batch(got, want)
    info = ctx.last_sort_info()
    assert info["key_bits"] <= 25, info                       # 30 s of spans + 2 x 2 s, not 400 days
    # events exactly one window before the first / after the last span still match; one nanosecond further they do not
    sp = w["spans"][:2000].copy()
    t_lo, t_hi = int(sp["ts_ns"].min()), int(sp["ts_ns"].max())
    W = 2_000_000_000
    edge = w["events"][:4000].copy()
    first, last = int(np.argmin(sp["ts_ns"])), int(np.argmax(sp["ts_ns"]))
    for k, (src, t) in enumerate(((first, t_lo - W), (first, t_lo - W - 1), (last, t_hi + W), (last, t_hi + W + 1))):
        edge["ts_ns"][k] = t
        edge["ident"][k] = sp["ident"][src]
        edge["trace_sig"][k] = (int(sp["trace"][src]) << 4) | 0
    got = ctx.process_batch(sp, edge, w["idents"])
    want = O.process_batch(sp, edge, w["idents"], indexed=False, nthreads=8)
    _eq_batch(got, want)
    # no span can match at all: nothing is kept, every supported pair is unmatched
    dead = sp.copy()
    dead["ts_ns"][:] = A.TS_ZERO
    got = ctx.process_batch(dead, edge, w["idents"])
    want = O.process_batch(dead, edge, w["idents"], indexed=False, nthreads=8)
    _eq_batch(got, want)
    assert (got["cand_event"] == A.NONE).all()


def test_brute_force_subsample(ctx):
    """the literal O(S x E) restatement of the reference loop on a sub-sample"""
    cfg = synth.config(50_000, 2_000, 16, 8, 2, seed=5, horizon_ns=10_000_000_000)
    w = synth.generate_host(cfg)
    got = ctx.process_batch(w["spans"], w["events"], w["idents"])
    want = O.process_batch(w["spans"], w["events"], w["idents"], indexed=False, nthreads=8)
    _eq_batch(got, want)


def test_edge_cases(ctx):
    cfg = synth.config(3000, 300, 8, 4, 2, seed=9, horizon_ns=5_000_000_000)
    w = synth.generate_host(cfg)
    ev, sp, ids = w["events"].copy(), w["spans"].copy(), w["idents"].copy()
    # empty sides
    for e, s in ((ev[:0], sp), (ev, sp[:0]), (ev[:0], sp[:0])):
        _eq_batch(ctx.process_batch(s, e, ids), O.process_batch(s, e, ids, indexed=True))
    # zero times, group 0, all events unsupported, base attributes, a wide time range (64-bit keys)
    ev2, sp2, ids2 = ev.copy(), sp.copy(), ids.copy()
    ev2["ts_ns"][::7] = A.TS_ZERO
    sp2["ts_ns"][::11] = A.TS_ZERO
    ids2["grp"][3] = 0
    base = np.full((len(sp2), A.NSIG), np.nan)
    base[::3, 0] = 1e9
    base[1::3, 3] = -5.0
    # (group 0 is the contract's "matches nothing": only the group-aware indexed oracle models it)
    _eq_batch(ctx.process_batch(sp2, ev2, ids2, attr_base=base), O.process_batch(sp2, ev2, ids2, attr_base=base, indexed=True))
    ids2["grp"][3] = ids["grp"][3]
    _eq_batch(ctx.process_batch(sp2, ev2, ids2, attr_base=base), O.process_batch(sp2, ev2, ids2, attr_base=base, indexed=False))
    ev3 = ev.copy()
    ev3["trace_sig"] |= 15
    _eq_batch(ctx.process_batch(sp, ev3, ids), O.process_batch(sp, ev3, ids, indexed=False))
    ev4, sp4 = ev.copy(), sp.copy()
    ev4["ts_ns"][::2] += 400 * 86400 * 10**9      # 400 days apart: > 2^53 ns, needs 64-bit keys
    sp4["ts_ns"][::2] += 400 * 86400 * 10**9
    ev4["ts_ns"][5] = -(1 << 62)
    sp4["ts_ns"][7] = (1 << 62)
    _eq_batch(ctx.process_batch(sp4, ev4, ids), O.process_batch(sp4, ev4, ids, indexed=False))
    # one group, every timestamp equal: a single key
    ev5, sp5 = ev.copy(), sp.copy()
    ev5["ts_ns"][:] = 12345
    sp5["ts_ns"][:] = 12345
    ids5 = ids.copy()
    ids5["grp"][:] = 1
    _eq_batch(ctx.process_batch(sp5, ev5, ids5), O.process_batch(sp5, ev5, ids5, indexed=False))
    # errors: ident out of range, oversized fanout / window
    bad = ev.copy()
    bad["ident"][10] = 9999
    import llm_slo_ebpf_toolkit_b200 as pkg
    with pytest.raises(pkg.LsgError) as e:
        ctx.process_batch(sp, bad, ids)
    assert e.value.code == -4
    with pytest.raises(pkg.LsgError):
        ctx.process_batch(sp, ev, ids, fanout=33)
    with pytest.raises(pkg.LsgError):
        ctx.process_batch(sp, ev, ids, window_ns=1 << 53)


def test_spans_sharing_a_time_quantum(ctx):
    """Spans are processed in (group, quantised time) order, so inside one quantum their exact times are not
    ordered; every one of them must still see the events at the very edge of its own +-W window."""
    W = 2_000_000_000
    base = synth.T0_NS
    n_s = 600
    rng = np.random.default_rng(5)
    sp = np.zeros(n_s, A.SPAN_DTYPE)
    sp["ts_ns"] = base + 10 * W + np.sort(rng.integers(0, 1 << 24, n_s))[::-1]    # decreasing inside quanta
    sp["ts_ns"][::7] += 1 << 25
    sp["ident"], sp["pid"], sp["conn"], sp["trace"], sp["incident"] = 0, 7, 3, np.arange(1, n_s + 1), A.NONE
    ev = np.zeros(4 * n_s + 5000, A.EVENT_DTYPE)
    k = np.arange(n_s)
    for r, dt in enumerate((W, -W, W - 1, -(W - 1))):                              # both edges, inclusive
        ev["ts_ns"][r * n_s + k] = sp["ts_ns"] + dt
        ev["trace_sig"][r * n_s + k] = (sp["trace"] << 4) | (r % 10)
    fill = slice(4 * n_s, None)                                                    # filler so that tiles have to be cut
    ev["ts_ns"][fill] = base + rng.integers(0, 20 * W, 5000)
    ev["trace_sig"][fill] = 2
    ev["pid"][fill] = 7
    ev["ident"], ev["value"] = 0, rng.random(len(ev))
    ids = np.array([(1, 1, 1, 1)], A.IDENT_DTYPE)
    got = ctx.process_batch(sp, ev, ids, W, 0.7, 8)
    want = O.process_batch(sp, ev, ids, W, 0.7, 8, indexed=False, nthreads=8)
    _eq_batch(got, want)
    assert (got["n_cand"] >= 4).all()


def test_match_pairs_vs_oracle(ctx):
    cfg = synth.config(100_000, 20_000, 16, 8, 2, seed=3, horizon_ns=40_000_000_000)
    w = synth.generate_host(cfg)
    ev, sp, ids = w["events"], w["spans"], w["idents"]
    # pair every event with a plausible span: the owner of its trace id, else the span of its join group nearest in time
    partner = np.zeros(len(ev), np.int64)
    sgrp, egrp = ids["grp"][sp["ident"]], ids["grp"][ev["ident"]]
    for g in np.unique(egrp):
        sidx = np.nonzero(sgrp == g)[0]
        sidx = sidx[np.argsort(sp["ts_ns"][sidx], kind="stable")]
        eidx = np.nonzero(egrp == g)[0]
        pos = np.clip(np.searchsorted(sp["ts_ns"][sidx], ev["ts_ns"][eidx]), 0, len(sidx) - 1)
        partner[eidx] = sidx[pos]
    tr = ev["trace_sig"] >> 4
    partner[tr != 0] = tr[tr != 0].astype(np.int64) - 1
    spans = sp[partner]
    rng = np.random.default_rng(1)
    em = rng.integers(0, 2, len(ev)).astype(np.uint8)
    et = rng.integers(0, 5, len(ev)).astype(np.uint8)
    for window, thr in ((0, 0.0), (300_000_000, 0.85), (50_000_000, 0.5), (5_000_000_000, 1.0)):
        t1, p1, e1 = ctx.match_pairs(spans, ev, ids, window, thr, em, et)
        t2, p2, e2 = O.match_pairs(spans, ev, ids, window, thr, em, et)
        assert np.array_equal(t1, t2) and np.array_equal(p1, p2) and bytes(e1) == bytes(e2)
    assert sorted(set(t2)) == [0, 1, 2, 3, 4] and min(e2.matched_by_tier) > 50


def test_attribute_features_confusion_vs_oracle(ctx):
    rng = np.random.default_rng(7)
    n = 50_000
    sig = rng.choice([np.nan, 0.0, 0.5, 1.0, 2.0, 5.0, 10.0, 39.999, 40.0, 60.0, 80.0, 250.0], size=(n, A.NSIG))
    sig[::97] = np.nan
    rule = rng.integers(0, 8, n).astype(np.uint8)
    post, top1, elev = ctx.attribute(sig, None, rule)
    opost, otop1, oelev = O.attribute(O.default_model(), sig, None, rule)
    assert np.array_equal(top1, otop1) and np.array_equal(elev, oelev)
    assert np.array_equal(np.isnan(post), np.isnan(opost))
    ok = ~np.isnan(opost)
    assert np.all(np.abs(post[ok] - opost[ok]) <= 1e-6 * opost[ok])     # north_star tolerance
    assert np.max(np.abs(post[ok] - opost[ok]) / opost[ok]) < 1e-12    # in practice a few ulp
    has = rng.integers(0, 2, n).astype(np.uint8)
    a1, a2 = ctx.attribute(sig, has, rule), O.attribute(O.default_model(), sig, has, rule)
    assert np.array_equal(a1[1], a2[1]) and np.array_equal(a1[2], a2[2])
    # confusion
    actual = rng.integers(0, 10, n).astype(np.uint8)
    exp_n = rng.integers(1, 4, n).astype(np.uint8)
    exp = rng.integers(0, 1 << 12, n).astype(np.uint32) & np.uint32(0x777)
    exp[::5] |= 0xF
    pred = np.where(top1 == 0xFF, 7, top1).astype(np.uint8)
    for thr, floor in ((0.10, 0.01), (0.0, 0.0), (0.5, 0.01)):
        c1 = ctx.confusion(actual, pred, exp, exp_n, post, thr, floor)
        c2 = O.confusion(actual, pred, exp, exp_n, opost, thr, floor)
        assert bytes(c1) == bytes(c2)
    assert bytes(ctx.confusion(actual, pred, exp, exp_n, None, 0.1, 0.01)) == bytes(O.confusion(actual, pred, exp, exp_n, None, 0.1, 0.01))
    # features
    ns, ni = 120_000, 9_000
    attr = rng.choice([np.nan, np.nan, -1.0, 0.0, 0.3, 7.0, 100.0, 5000.0], size=(ns, A.NSIG))
    inc = np.sort(rng.integers(0, ni, ns)).astype(np.uint32)
    inc[::13] = A.NONE
    inc[5::1000] = rng.integers(0, ni, len(inc[5::1000]))
    s1, f1 = ctx.incident_features(attr, inc, ni, want_features=True)
    s2, f2 = O.incident_features(attr, inc, ni, want_features=True)
    assert np.array_equal(np.nan_to_num(s1, nan=-1e300), np.nan_to_num(s2, nan=-1e300))
    assert np.array_equal(f1["count"], f2["count"]) and np.array_equal(f1["sketch"], f2["sketch"])
    assert np.allclose(f1["sum"], f2["sum"], rtol=1e-6, atol=1e-9)


def _dev_setup(ctx, cfg, **kw):
    from llm_slo_ebpf_toolkit_b200.device import DeviceWorkload
    return DeviceWorkload(ctx, cfg, **kw)


def test_device_generator_equals_host_generator(ctx):
    for multi in (0, 1):
        cfg = synth.config(100_000, 10_000, 32, 8, 2, seed=11, multi_fault=multi, rank=multi, world=2)
        w = synth.generate_host(cfg)
        d = _dev_setup(ctx, cfg)
        assert d.host("events", A.EVENT_DTYPE).tobytes() == w["events"].tobytes()
        assert d.host("spans", A.SPAN_DTYPE).tobytes() == w["spans"].tobytes()
        assert d.host("idents", A.IDENT_DTYPE).tobytes() == w["idents"].tobytes()
        assert np.array_equal(d.host("inc_actual", np.uint8)[: len(w["inc_actual"])], w["inc_actual"])
        assert np.array_equal(d.host("inc_expected", np.uint32)[: len(w["inc_actual"])], w["inc_expected"])


def test_sort_events_exact(ctx):
    import torch
    cfg = synth.config(300_000, 1_000, 32, 8, 2, seed=13)
    w = synth.generate_host(cfg)
    ev = w["events"].copy()
    ev["ts_ns"][::50] = ev["ts_ns"][1::50][: len(ev["ts_ns"][::50])]   # duplicate timestamps: stability is visible
    ev["ts_ns"][::301] = A.TS_ZERO
    d = _dev_setup(ctx, cfg, generate=False)
    w2 = dict(w)
    w2["events"] = ev
    d.upload(w2)
    n = len(ev)
    out = torch.empty(n * 32, dtype=torch.uint8, device="cuda")
    perm = torch.empty(n, dtype=torch.int32, device="cuda")
    keys = torch.empty(n, dtype=torch.int64, device="cuda")
    ctx.call_dev("lsg_sort_events_dev", d.events.data_ptr(), n, d.idents.data_ptr(), d.n_idents, out.data_ptr(),
                 perm.data_ptr(), keys.data_ptr())
    perm = perm.cpu().numpy().view(np.uint32)
    keys = keys.cpu().numpy().view(np.uint64)
    rec = out.cpu().numpy().view(A.EVENT_DTYPE)
    assert np.array_equal(np.sort(perm), np.arange(n, dtype=np.uint32))           # a permutation
    assert rec.tobytes() == ev[perm].tobytes()                                     # records follow it
    assert np.all(keys[:-1] <= keys[1:])                                           # sorted
    same = keys[:-1] == keys[1:]
    assert np.all(perm[:-1][same] < perm[1:][same])                                # stable
    grp = w["idents"]["grp"][ev["ident"]]
    live = ((ev["trace_sig"] & 15) < A.NSIG) & (ev["ts_ns"] != A.TS_ZERO) & (grp != 0)
    n_live = int(live.sum())
    assert live[perm[:n_live]].all() and not live[perm[n_live:]].any()             # dead records sort last
    order = np.lexsort((np.arange(n), ev["ts_ns"], grp))
    assert np.array_equal(perm[:n_live], order[live[order]])                       # = stable sort by (grp, ts)


def _oracle_pipeline(w, window_ns=0, thr=0.0, fanout=3, cov=0.10):
    j = O.process_batch(w["spans"], w["events"], w["idents"], window_ns, thr, fanout, indexed=True, nthreads=8)
    ni = len(w["inc_actual"])
    sig = O.incident_features(j["attr"], w["spans"]["incident"], ni)
    post, top1, elev = O.attribute(O.default_model(), sig)
    conf = O.confusion(w["inc_actual"], top1, w["inc_expected"], w["inc_expected_n"], post, cov, 0.01)
    return j, sig, post, top1, conf


@pytest.mark.parametrize("multi", [0, 1])
def test_pipeline_dev_vs_oracle(ctx, multi):
    import torch
    cfg = synth.config(1_000_000, 100_000, 64, 8, 2, seed=42, multi_fault=multi, horizon_ns=360_000_000_000)
    w = synth.generate_host(cfg)
    d = _dev_setup(ctx, cfg)
    d.step()
    d.step()        # a second pass over the same buffers gives the same result
    torch.cuda.synchronize()
    j, sig, post, top1, conf = _oracle_pipeline(w)
    ns, ni = d.n_spans, d.n_incidents
    assert np.array_equal(d.host("cand_event", np.uint32, (ns, 3)), j["cand_event"])
    assert np.array_equal(d.host("cand_tier", np.uint8)[: ns * 3].reshape(ns, 3), j["cand_tier"])
    assert np.array_equal(np.nan_to_num(d.host("attr", np.float64, (ns, A.NSIG)), nan=-1e300), np.nan_to_num(j["attr"], nan=-1e300))
    assert np.array_equal(d.host("conf", np.float64)[:ns], j["conf"])
    assert np.array_equal(np.nan_to_num(d.host("inc_signals", np.float64, (ni, A.NSIG)), nan=-1e300), np.nan_to_num(sig, nan=-1e300))
    assert np.array_equal(d.host("top1", np.uint8)[:ni], top1)
    gp = d.host("posterior", np.float64, (ni, A.NDOM))
    ok = ~np.isnan(post)
    assert np.array_equal(np.isnan(gp), np.isnan(post)) and np.all(np.abs(gp[ok] - post[ok]) <= 1e-6 * post[ok])
    c = d.read_counters()
    assert bytes(c.confusion) == bytes(conf)
    assert c.debug.as_tuple() == j["debug"].as_tuple()
    assert (c.events, c.spans, c.incidents) == (cfg.n_events, cfg.n_spans, ni)
    assert c.candidates == int(j["n_cand"].sum()) and c.kept == int((j["cand_event"] != A.NONE).sum())


def test_full_c2_properties(ctx):
    """BASELINE.json configs[1] at full size (10 M events x 1 M spans): invariants that do not need the
    brute force, plus exact comparison of sampled spans against the oracle on their own join group."""
    import torch
    cfg = synth.c2()
    d = _dev_setup(ctx, cfg)
    d.step()
    torch.cuda.synchronize()
    c = d.read_counters()
    ns, ni = d.n_spans, d.n_incidents
    m = np.array(c.confusion.matrix)
    assert m.sum() + 0 == c.confusion.n == ni                                   # every incident lands in one cell
    assert c.confusion.correct == np.trace(m[:8]) and c.confusion.partial >= c.confusion.correct
    assert np.array(c.confusion.coverage).sum() == ni
    cand = d.host("cand_event", np.uint32, (ns, 3))
    tier = d.host("cand_tier", np.uint8)[: ns * 3].reshape(ns, 3)
    kept = cand != A.NONE
    assert c.kept == kept.sum() and c.candidates >= c.kept and c.debug.fanout_dropped == c.candidates - c.kept
    assert np.all(kept[:, :-1] >= kept[:, 1:])                                  # kept candidates are a prefix
    assert np.all((tier[:, :-1] <= tier[:, 1:]) | ~kept[:, 1:])                 # confidence never increases along it
    ev = d.host("events", A.EVENT_DTYPE)
    unsup = int(((ev["trace_sig"] & 15) >= A.NSIG).sum())
    assert c.debug.unsupported_type == unsup * ns
    assert c.debug.unmatched + c.debug.low_confidence + c.candidates == (cfg.n_events - unsup) * ns
    # the whole batch against the indexed CPU oracle (a few seconds on the host cores)
    w = synth.generate_host(cfg)
    j = O.process_batch(w["spans"], w["events"], w["idents"], indexed=True, nthreads=O.hw_threads())
    assert c.debug.as_tuple() == j["debug"].as_tuple()
    assert c.candidates == int(j["n_cand"].sum())
    assert np.array_equal(cand, j["cand_event"]) and np.array_equal(tier, j["cand_tier"])
    # ... and everything downstream of the join at full size: enriched attributes, per-incident maxima, labels,
    # posteriors (1e-6), the confusion block
    attr_full = d.host("attr", np.float64, (ns, A.NSIG))
    assert np.array_equal(attr_full, j["attr"], equal_nan=True)
    assert np.array_equal(d.host("conf", np.float64)[:ns], j["conf"])
    osig = O.incident_features(j["attr"], w["spans"]["incident"], ni)
    assert np.array_equal(d.host("inc_signals", np.float64, (ni, A.NSIG)), osig, equal_nan=True)
    opost, otop1, _ = O.attribute(O.default_model(), osig)
    assert np.array_equal(d.host("top1", np.uint8)[:ni], otop1)
    assert np.allclose(d.host("posterior", np.float64, (ni, A.NDOM)), opost, rtol=1e-6, atol=0, equal_nan=True)
    oconf = O.confusion(w["inc_actual"], otop1, w["inc_expected"], w["inc_expected_n"], opost, 0.10, 0.01)
    assert bytes(c.confusion) == bytes(oconf)
    # sampled join groups, exact
    sp, ids = d.host("spans", A.SPAN_DTYPE), d.host("idents", A.IDENT_DTYPE)
    attr = d.host("attr", np.float64, (ns, A.NSIG))
    for g in (1, 17, 32):
        smask = ids["grp"][sp["ident"]] == g
        emask = ids["grp"][ev["ident"]] == g
        sidx, eidx = np.nonzero(smask)[0][:4000], np.nonzero(emask)[0]
        want = O.process_batch(sp[sidx], ev[eidx], ids, indexed=True, nthreads=8)
        wc = np.where(want["cand_event"] == A.NONE, A.NONE, eidx[np.minimum(want["cand_event"], len(eidx) - 1)]).astype(np.uint32)
        assert np.array_equal(cand[sidx], wc)
        assert np.array_equal(tier[sidx], want["cand_tier"])
        assert np.array_equal(np.nan_to_num(attr[sidx], nan=-1e300), np.nan_to_num(want["attr"], nan=-1e300))


def test_host_pipeline_blocking_and_double_buffered(ctx):
    """lsg_pipeline (host buffers, blocking) and lsg_pipeline_submit / _wait (two batches in flight) give the device
    pipeline's results, batch after batch."""
    import torch
    from llm_slo_ebpf_toolkit_b200.device import DeviceWorkload
    batches = []
    for seed in (1, 2, 3):
        cfg = synth.config(300_000, 30_000, 32, 8, 2, seed=seed, horizon_ns=100_000_000_000, multi_fault=seed & 1)
        w = synth.generate_host(cfg)
        d = DeviceWorkload(ctx, cfg)
        d.step()
        torch.cuda.synchronize()
        batches.append((cfg, w, d.read_counters(), d.host("cand_event", np.uint32).copy(), d.host("posterior", np.float64).copy()))
    f = 3

    def host_args(w):
        ns, ni = len(w["spans"]), len(w["inc_actual"])
        keep = {k: np.ascontiguousarray(w[k]) for k in ("spans", "events", "idents", "inc_actual", "inc_expected", "inc_expected_n")}
        outs = {"ce": np.zeros(ns * f, np.uint32), "ct": np.zeros(ns * f, np.uint8), "attr": np.zeros(ns * A.NSIG), "conf": np.zeros(ns),
                "sig": np.zeros(ni * A.NSIG), "post": np.zeros(ni * A.NDOM), "top1": np.zeros(ni, np.uint8)}
        args = A.PipelineArgs(keep["spans"].ctypes.data, ns, keep["events"].ctypes.data, len(w["events"]), keep["idents"].ctypes.data,
                              len(w["idents"]), ni, keep["inc_actual"].ctypes.data, keep["inc_expected"].ctypes.data,
                              keep["inc_expected_n"].ctypes.data, 0, 0.0, f, 0.10, outs["ce"].ctypes.data, outs["ct"].ctypes.data,
                              outs["attr"].ctypes.data, outs["conf"].ctypes.data, outs["sig"].ctypes.data, outs["post"].ctypes.data,
                              outs["top1"].ctypes.data, None)
        return args, keep, outs

    def check(c, outs, ref):
        _cfg, _w, rc, rce, rpost = ref
        assert bytes(c) == bytes(rc)
        assert np.array_equal(outs["ce"], rce[: len(outs["ce"])])
        assert np.array_equal(np.nan_to_num(outs["post"], nan=-1.0), np.nan_to_num(rpost[: len(outs["post"])], nan=-1.0))

    for ref in batches:                                   # blocking form
        args, keep, outs = host_args(ref[1])
        c = A.Counters()
        ctx.call_dev("lsg_pipeline", C.byref(args), C.byref(c))
        check(c, outs, ref)
    inflight = {}
    order = [0, 1, 2, 1, 0, 2]                            # double-buffered form: six batches through two slots
    for i, b in enumerate(order):
        slot = i & 1
        if slot in inflight:
            c = A.Counters()
            ctx.call_dev("lsg_pipeline_wait", slot, C.byref(c))
            check(c, inflight[slot][2], batches[inflight[slot][0]])
        args, keep, outs = host_args(batches[b][1])
        ctx.call_dev("lsg_pipeline_submit", C.byref(args), slot)
        inflight[slot] = (b, keep, outs, args)
    for slot in (0, 1):
        c = A.Counters()
        ctx.call_dev("lsg_pipeline_wait", slot, C.byref(c))
        check(c, inflight[slot][2], batches[inflight[slot][0]])
    import llm_slo_ebpf_toolkit_b200 as pkg
    with pytest.raises(pkg.LsgError):
        ctx.call_dev("lsg_pipeline_wait", 0, None)        # nothing in flight
    # a span whose ident is outside the identity table fails the whole call (LSG_ERR_RANGE), in the pipeline too
    w_bad = dict(batches[0][1])
    w_bad["spans"] = w_bad["spans"].copy()
    w_bad["spans"]["ident"][17] = 1 << 30
    args, keep, outs = host_args(w_bad)
    with pytest.raises(pkg.LsgError) as e:
        ctx.call_dev("lsg_pipeline", C.byref(args), C.byref(A.Counters()))
    assert e.value.code == -4
    with pytest.raises(pkg.LsgError) as e:
        ctx.process_batch(w_bad["spans"], w_bad["events"], w_bad["idents"])
    assert e.value.code == -4
    w_bad = dict(batches[0][1])                           # the same for an event
    w_bad["events"] = w_bad["events"].copy()
    w_bad["events"]["ident"][5] = 1 << 30
    args, keep, outs = host_args(w_bad)
    with pytest.raises(pkg.LsgError) as e:
        ctx.call_dev("lsg_pipeline", C.byref(args), C.byref(A.Counters()))
    assert e.value.code == -4
    ctx.call_dev("lsg_pipeline_submit", C.byref(args), 0)
    with pytest.raises(pkg.LsgError) as e:
        ctx.call_dev("lsg_pipeline_wait", 0, C.byref(A.Counters()))
    assert e.value.code == -4
    args, keep, outs = host_args(batches[1][1])           # and the context is usable afterwards
    c = A.Counters()
    ctx.call_dev("lsg_pipeline", C.byref(args), C.byref(c))
    check(c, outs, batches[1])


def test_compact_span_outputs_rebuild_the_dense_ones(ctx):
    """lsg_pipeline_args.cand_value / cand_sig: the kept candidates' value and signal are enough to rebuild a span's
    attribute row (max per signal, correlator.go:106-112) and confidence (first candidate's tier, correlator.go:113-118),
    so a host caller may leave attr_out / conf_out NULL and download a third of the bytes."""
    for fanout, ne, ns, horizon in ((3, 200_000, 20_000, 60), (8, 60_000, 6_000, 10), (12, 30_000, 300, 1)):
        cfg = synth.config(ne, ns, 16, 8, 2, seed=40 + fanout, horizon_ns=horizon * 1_000_000_000)
        w = synth.generate_host(cfg)
        n_s, n_i = len(w["spans"]), len(w["inc_actual"])
        keep = {k: np.ascontiguousarray(w[k]) for k in ("spans", "events", "idents", "inc_actual", "inc_expected", "inc_expected_n")}

        def run(dense, compact):
            o = {"ce": np.zeros(n_s * fanout, np.uint32), "ct": np.zeros(n_s * fanout, np.uint8),
                 "attr": np.zeros(n_s * A.NSIG), "conf": np.zeros(n_s), "sig": np.zeros(n_i * A.NSIG),
                 "post": np.zeros(n_i * A.NDOM), "top1": np.zeros(n_i, np.uint8),
                 "cv": np.full(n_s * fanout, -1.0), "cs": np.full(n_s * fanout, 7, np.uint8)}
            args = A.PipelineArgs(keep["spans"].ctypes.data, n_s, keep["events"].ctypes.data, len(w["events"]),
                                  keep["idents"].ctypes.data, len(w["idents"]), n_i, keep["inc_actual"].ctypes.data,
                                  keep["inc_expected"].ctypes.data, keep["inc_expected_n"].ctypes.data, 0, 0.0, fanout, 0.10,
                                  o["ce"].ctypes.data, o["ct"].ctypes.data, o["attr"].ctypes.data if dense else None,
                                  o["conf"].ctypes.data if dense else None, o["sig"].ctypes.data, o["post"].ctypes.data,
                                  o["top1"].ctypes.data, None, o["cv"].ctypes.data if compact else None,
                                  o["cs"].ctypes.data if compact else None)
            c = A.Counters()
            ctx.call_dev("lsg_pipeline", C.byref(args), C.byref(c))
            return o, c

        dense, c_dense = run(True, False)
        both, c_both = run(True, True)
        comp, c_comp = run(False, True)
        assert bytes(c_dense) == bytes(c_both) == bytes(c_comp)
        for k in ("ce", "ct", "sig", "post", "top1"):
            assert np.array_equal(dense[k], comp[k], equal_nan=True) and np.array_equal(dense[k], both[k], equal_nan=True), k
        assert np.array_equal(both["cv"], comp["cv"]) and np.array_equal(both["cs"], comp["cs"])
        assert not comp["attr"].any() and not comp["conf"].any()          # not requested: left untouched
        ce, ct = comp["ce"].reshape(n_s, fanout), comp["ct"].reshape(n_s, fanout)
        cv, cs = comp["cv"].reshape(n_s, fanout), comp["cs"].reshape(n_s, fanout)
        has = ce != A.NONE
        assert ((cs == 0xFF) == ~has).all() and (cv[~has] == 0).all()
        assert np.array_equal(cs[has], (w["events"]["trace_sig"][ce[has]] & 15).astype(np.uint8))
        assert np.array_equal(cv[has], w["events"]["value"][ce[has]])
        attr = np.full((n_s, A.NSIG), np.nan)
        for r in range(fanout):                                            # max per signal over the kept candidates
            rows = np.nonzero(has[:, r])[0]
            cur = attr[rows, cs[rows, r]]
            attr[rows, cs[rows, r]] = np.where(np.isnan(cur) | (cv[rows, r] > cur), cv[rows, r], cur)
        assert np.array_equal(attr.reshape(-1), dense["attr"], equal_nan=True)
        conf = np.where(has[:, 0], np.array(A.TIER_CONF)[ct[:, 0]], 0.0)
        assert np.array_equal(conf, dense["conf"])


@pytest.mark.parametrize("case", range(24))
def test_randomized_configs_vs_indexed_oracle(ctx, case):
    """Seeded random draws of (sizes, pods, class shape, horizon, window, threshold, fanout): every special path of the
    join (counted service+node tier, note-list overflow, spans beyond a tile, fanout above the register list, radix keys)
    is reached by some of them; each must equal the indexed oracle field for field."""
    rng = np.random.default_rng(1000 + case)
    n_pods = int(rng.choice([2, 4, 8, 16, 64, 256]))
    ppg = int(rng.choice([1, 2, 4]))
    ppg = min(ppg, n_pods)
    ppn = max(ppg, int(rng.choice([ppg, 2 * ppg, 4 * ppg])))
    ppn = min(ppn, n_pods)
    n_spans = int(rng.integers(200, 20_000))
    n_events = int(n_spans * rng.choice([2, 5, 10, 20]))
    horizon_s = float(rng.choice([0.5, 5, 60, 3600, 400 * 86400]))
    window_ms = int(rng.choice([1, 40, 100, 249, 250, 251, 499, 500, 501, 2000, 10_000, 120_000]))
    thr = float(rng.choice([0.5, 0.65, 0.66, 0.7, 0.8, 0.81, 0.9, 0.95, 1.0, 1.01]))
    fanout = int(rng.choice([1, 2, 3, 4, 8, 9, 32]))
    cfg = synth.config(n_events, n_spans, n_pods, ppn, ppg, seed=77 + case, horizon_ns=int(horizon_s * 1e9),
                       multi_fault=case & 1)
    w = synth.generate_host(cfg)
    ids = w["idents"].copy()
    if case % 5 == 4:
        ids["service"][rng.integers(0, len(ids), max(1, len(ids) // 4))] = 0   # rows without a (service, node) class
    got = ctx.process_batch(w["spans"], w["events"], ids, window_ms * 1_000_000, thr, fanout)
    want = O.process_batch(w["spans"], w["events"], ids, window_ms * 1_000_000, thr, fanout, indexed=True, nthreads=8)
    _eq_batch(got, want)


def test_stated_key_range_skips_the_readback_and_is_checked(ctx):
    """lsg_pipeline_args.key_hint: the caller states the spans' time range and the largest group; results are those
    of the un-hinted run, a statement that does not hold voids the batch (counters.bad_idents)."""
    import torch
    from llm_slo_ebpf_toolkit_b200.device import DeviceWorkload
    cfg = synth.config(300_000, 30_000, 32, 8, 2, seed=9, horizon_ns=100_000_000_000, multi_fault=1)
    plain = DeviceWorkload(ctx, cfg)
    plain.step()
    torch.cuda.synchronize()
    want = plain.read_counters()
    assert want.bad_idents == 0
    hinted = DeviceWorkload(ctx, cfg)
    hinted.state_key_range()
    for _ in range(3):                       # back to back without a host synchronisation in between
        hinted.step()
    torch.cuda.synchronize()
    got = hinted.read_counters()
    assert bytes(got) == bytes(want)
    for name, dt in (("cand_event", np.uint32), ("cand_tier", np.uint8), ("top1", np.uint8)):
        assert np.array_equal(hinted.host(name, dt), plain.host(name, dt)), name
    assert np.array_equal(hinted.host("attr", np.float64), plain.host("attr", np.float64), equal_nan=True)
    assert np.array_equal(hinted.host("posterior", np.float64), plain.host("posterior", np.float64), equal_nan=True)
    # a wider statement is still true: same results
    hinted.state_key_range(span_ts_lo=cfg.t0_ns - 10**12, span_ts_hi=cfg.t0_ns + cfg.horizon_ns + 10**12, max_grp=1000)
    hinted.step()
    torch.cuda.synchronize()
    assert bytes(hinted.read_counters()) == bytes(want)
    # statements that do not hold: a time range that ends too early, a group bound that is too small
    hinted.state_key_range(span_ts_hi=cfg.t0_ns + cfg.horizon_ns // 2)
    hinted.step()
    torch.cuda.synchronize()
    assert hinted.read_counters().bad_idents != 0
    hinted.state_key_range(max_grp=3)
    hinted.step()
    torch.cuda.synchronize()
    assert hinted.read_counters().bad_idents != 0
    hinted.state_key_range()
    hinted.step()
    torch.cuda.synchronize()
    assert bytes(hinted.read_counters()) == bytes(want)
    # with the statement nothing is validated up front: bad idents must stay in bounds and surface in the counters
    w = synth.generate_host(cfg)
    for field in ("spans", "events"):
        bad = dict(w)
        bad[field] = w[field].copy()
        bad[field]["ident"][123] = 0x7FFFFFFF
        hinted.upload(bad)
        hinted.step()
        torch.cuda.synchronize()
        assert hinted.read_counters().bad_idents != 0, field
    hinted.upload(w)
    hinted.step()
    torch.cuda.synchronize()
    assert bytes(hinted.read_counters()) == bytes(want)
