"""Proves the packed-record C++ oracle (oracle/oracle.cpp) equal to the string-keyed restatement
(oracle/ref_str.py) through the interner, on the reference fixtures and on randomised batches; and
the indexed variant equal to the brute force.  CPU only."""
import math

import numpy as np
import pytest

from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200.interner import Interner
from oracle import cpu as O
from oracle import ref_str as R
from tests.fixtures import load_golden, pairs_from_records, samples_from_records
from tests.randgen import random_batch

TIER_CODE = {n: i for i, n in enumerate(A.TIER_NAMES)}


def test_pairs_fixture_packed_equals_string():
    pairs = pairs_from_records(load_golden("labeled_pairs"))
    it = Interner()
    sp = it.span_records([p.Span for p in pairs])
    ev = it.event_records([p.Signal for p in pairs])
    em = np.array([p.ExpectedMatch for p in pairs], np.uint8)
    et = np.array([TIER_CODE[p.ExpectedTier] for p in pairs], np.uint8)
    for window_ms, thr in [(2000, 0.7), (50, 0.7), (300, 0.85), (5000, 0.5), (2000, 1.0), (0, 0.0)]:
        tier, pred, e = O.match_pairs(sp, ev, it.ident_table(), window_ms * R.MS, thr, em, et)
        rep, preds = R.EvaluateLabeledPairs(pairs, window_ms * R.MS, thr)
        assert [TIER_CODE[p.Tier] for p in preds] == list(tier)
        assert [p.Predicted for p in preds] == [bool(x) for x in pred]
        assert (e.tp, e.fp, e.fn, e.tn) == (rep.TruePositive, rep.FalsePositive, rep.FalseNegative, rep.TrueNegative)
        assert (e.tier_correct, e.tier_comparable) == (rep._tier_correct, rep._tier_comparable)
    tier, pred, e = O.match_pairs(sp, ev, it.ident_table(), 2000 * R.MS, 0.7, em, et)
    assert (e.tp, e.fp, e.fn, e.tn) == (40, 0, 0, 15)
    assert list(e.matched_by_tier) == [6, 17, 14, 9, 9]


def _compare_batch(spans, signals, corr, indexed, nthreads=1):
    it = Interner()
    sp = it.span_records([R.SpanRef(s.TraceID, s.Service, s.Node, s.Pod, s.PID, s.ConnTuple, s.Timestamp) for s in spans])
    ev = it.event_records(signals)
    base = np.full((len(spans), A.NSIG), np.nan)
    for i, s in enumerate(spans):
        for k, v in (s.Attributes or {}).items():
            if k in R.ATTR.values():
                base[i, list(R.ATTR.values()).index(k)] = v
    got = O.process_batch(sp, ev, it.ident_table(), corr.WindowMS * R.MS, corr.EnrichmentThreshold, corr.MaxJoinFanout,
                          base, indexed=indexed, nthreads=nthreads)
    want = corr.ProcessBatch(spans, signals)
    d = got["debug"]
    assert (d.unmatched, d.low_confidence, d.fanout_dropped, d.unsupported_type) == \
        (want.Debug.Unmatched, want.Debug.LowConfidence, want.Debug.FanoutDropped, want.Debug.UnsupportedType)
    keys = list(R.ATTR.values())
    fan = corr.MaxJoinFanout if corr.MaxJoinFanout > 0 else 3
    for i, w in enumerate(want.Spans):
        cands = [c.Index for c in w._candidates]
        assert [int(x) for x in got["cand_event"][i] if x != A.NONE] == cands, i
        assert [int(x) for x in got["cand_tier"][i][:len(cands)]] == [TIER_CODE[c.Decision.Tier] for c in w._candidates]
        for j, k in enumerate(keys):
            if k in w.Attributes:
                assert got["attr"][i, j] == w.Attributes[k]
            else:
                assert math.isnan(got["attr"][i, j])
        assert got["conf"][i] == w.Attributes.get(R.ATTR_CORRELATION_CONF, 0.0)
        assert got["retrieval"][i] == w.Attributes.get(R.ATTR_RETRIEVAL_KERNEL_MS, 0.0) or \
            (got["retrieval"][i] <= 0 and R.ATTR_RETRIEVAL_KERNEL_MS not in w.Attributes)
    return got


@pytest.mark.parametrize("indexed", [False, True])
@pytest.mark.parametrize("seed", range(6))
def test_process_batch_packed_equals_string(seed, indexed):
    spans, signals = random_batch(seed, 24, 160, n_pods=3 + seed % 3)
    corr = R.Correlator([2000, 2000, 300, 50, 5000, 0][seed], [0.7, 0.85, 0.7, 0.5, 0.95, 0.0][seed], [3, 1, 8, 3, 2, 0][seed])
    _compare_batch(spans, signals, corr, indexed, nthreads=1 + seed % 3)


def test_process_batch_edge_shapes():
    spans, signals = random_batch(7, 5, 40)
    for sp, sg in [([], signals), (spans, []), ([], [])]:
        _compare_batch(sp, sg, R.New(), False)
        _compare_batch(sp, sg, R.New(), True)


def test_indexed_equals_brute_larger():
    spans, signals = random_batch(11, 300, 4000, n_pods=6, dense=False)
    it = Interner()
    sp = it.span_records([R.SpanRef(s.TraceID, s.Service, s.Node, s.Pod, s.PID, s.ConnTuple, s.Timestamp) for s in spans])
    ev = it.event_records(signals)
    a = O.process_batch(sp, ev, it.ident_table(), indexed=False, nthreads=4)
    b = O.process_batch(sp, ev, it.ident_table(), indexed=True, nthreads=3)
    for k in ("cand_event", "cand_tier", "conf", "retrieval", "n_cand"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(np.nan_to_num(a["attr"], nan=-1e300), np.nan_to_num(b["attr"], nan=-1e300))
    assert a["debug"].as_tuple() == b["debug"].as_tuple()


def test_attribute_packed_equals_string():
    samples = samples_from_records(load_golden("multi_fault_samples"))
    ba = R.NewBayesianAttributor()
    want = [ba.AttributeSample(s) for s in samples]
    sig = np.full((len(samples), A.NSIG), np.nan)
    for i, s in enumerate(samples):
        for k, v in s.Signals.items():
            sig[i, A.SIGNAL_INDEX[k]] = v
    post, top1, elev = O.attribute(O.default_model(), sig)
    for i, w in enumerate(want):
        assert A.DOMAINS[top1[i]] == w.PredictedFaultDomain
        assert post[i, top1[i]] == w.Confidence          # same libm, same summation order: bit-equal
        for d, p, evidence in w.FaultHypotheses:
            assert post[i, A.DOMAIN_INDEX[d]] == p
        assert abs(post[i].sum() - 1.0) < 1e-12
    # elevated mask vs thresholds
    for i, s in enumerate(samples):
        want_mask = sum(1 << A.SIGNAL_INDEX[k] for k, v in s.Signals.items() if v >= R.SIGNAL_THRESHOLDS[k])
        assert elev[i] == want_mask


def test_confusion_packed_equals_string():
    samples = samples_from_records(load_golden("multi_fault_samples"))
    attr = R.BuildAttributionsBayes(samples, None)
    n = len(samples)
    actual = np.array([A.DOMAIN_INDEX[s.ExpectedDomain or R.MapFaultLabel(s.FaultLabel)] for s in samples], np.uint8)
    pred = np.array([A.DOMAIN_INDEX[a.PredictedFaultDomain] for a in attr], np.uint8)
    exp = np.zeros(n, np.uint32)
    exp_n = np.zeros(n, np.uint8)
    post = np.full((n, A.NDOM), np.nan)
    for i, (s, a) in enumerate(zip(samples, attr)):
        lst = s.ExpectedDomains or [s.ExpectedDomain or R.MapFaultLabel(s.FaultLabel)]
        exp[i] = sum(A.DOMAIN_INDEX.get(d, 15) << (4 * k) for k, d in enumerate(lst))
        exp_n[i] = len(lst)
        for d, p, _ in a.FaultHypotheses:
            post[i, A.DOMAIN_INDEX[d]] = p
    out = O.confusion(actual, pred, exp, exp_n, post, 0.10)
    assert out.n == 55 and out.correct == 29 and out.partial == 54
    cov = sum(out.coverage[f][ln] * f / ln for f in range(9) for ln in range(1, 9)) / out.n
    assert abs(cov - R.CoverageAccuracy(samples, attr, 0.10)) < 1e-12
    m = {(A.DOMAINS[a], A.DOMAINS[p]): out.matrix[a][p] for a in range(8) for p in range(8) if out.matrix[a][p]}
    assert m == R.BuildConfusionMatrix(samples, attr)
    # raw posteriors + the 0.01 floor give the same answer as the explicit hypothesis list
    sig = np.full((n, A.NSIG), np.nan)
    for i, s in enumerate(samples):
        for k, v in s.Signals.items():
            sig[i, A.SIGNAL_INDEX[k]] = v
    raw, _, _ = O.attribute(O.default_model(), sig)
    out2 = O.confusion(actual, pred, exp, exp_n, raw, 0.10, 0.01)
    assert bytes(out2) == bytes(out)


def test_incident_features_and_retry_storm():
    rng = np.random.default_rng(3)
    attr = rng.choice([np.nan, 0.0, 0.3, 7.0, 100.0, 5000.0], size=(200, A.NSIG))
    inc = rng.integers(0, 20, 200).astype(np.uint32)
    inc[::17] = A.NONE
    sig, feat = O.incident_features(attr, inc, 20, want_features=True)
    for k in range(20):
        rows = attr[inc == k]
        for s in range(A.NSIG):
            col = rows[:, s][~np.isnan(rows[:, s])]
            if len(col) == 0:
                assert math.isnan(sig[k, s]) and feat[k, s]["count"] == 0
            else:
                assert sig[k, s] == col.max() and feat[k, s]["count"] == len(col)
                assert abs(feat[k, s]["sum"] - col.sum()) < 1e-9
                assert feat[k, s]["sketch"].sum() == len(col)
    # retry storm replay vs the string detector
    det = R.RetryStormDetector(10 * R.SECOND, 5)
    pods = rng.integers(0, 3, 300).astype(np.uint32)
    ts = np.sort(rng.integers(0, 60_000, 300)).astype(np.int64) * R.MS
    cnt, storm = O.retry_storm(pods, ts, 3, 10 * R.SECOND, 5)
    for i in range(300):
        assert det.Record("p%d" % pods[i], int(ts[i])) == bool(storm[i])
        assert det.Count("p%d" % pods[i], int(ts[i])) == cnt[i]
