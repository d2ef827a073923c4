"""Pins oracle/ref_str.py (the string-keyed restatement) to the reference's fixtures and to the
assertions of the reference's own unit tests, transcribed case by case.

Sources (reference tree): pkg/correlation/dns_test.go, evaluator_test.go, retry_storm_test.go,
pkg/otel/processor/ebpfcorrelator/correlator_test.go, pkg/attribution/bayesian_test.go,
pipeline_test.go, mapper_test.go; golden values of SURVEY.md §4.
"""
import math

import pytest

from oracle import ref_str as R
from tests.fixtures import load_golden, pairs_from_records, samples_from_records

NOW = 1_771_372_800_000_000_000  # any fixed "time.Now()"
MS, S = R.MS, R.SECOND


# ---- pkg/correlation/dns_test.go:8-87 TestMatchTiers
@pytest.mark.parametrize("signal,conf,tier", [
    (R.SignalRef("dns_latency_ms", TraceID="trace-1", Timestamp=NOW + 1 * S), 1.0, "trace_id_exact"),
    (R.SignalRef("dns_latency_ms", Pod="pod-a", PID=42, Timestamp=NOW + 50 * MS), 0.9, "pod_pid_100ms"),
    (R.SignalRef("dns_latency_ms", Pod="pod-a", ConnTuple="10.0.0.1:1234->10.0.0.2:53/udp", Timestamp=NOW + 200 * MS),
     0.8, "pod_conn_250ms"),
    (R.SignalRef("dns_latency_ms", Service="chat", Node="node-a", Timestamp=NOW + 400 * MS), 0.65, "service_node_500ms"),
])
def test_match_tiers(signal, conf, tier):
    span = R.SpanRef("trace-1", "chat", "node-a", "pod-a", 42, "10.0.0.1:1234->10.0.0.2:53/udp", NOW)
    d = R.Match(span, signal, R.DEFAULT_WINDOW)
    assert d.Matched and d.Confidence == conf and d.Tier == tier


# ---- dns_test.go:89-132 TestEnrichDNSRespectsThreshold
def test_enrich_dns_respects_threshold():
    span = R.SpanRef(Service="chat", Node="node-a", Timestamp=NOW)
    sig = R.SignalRef("dns_latency_ms", Service="chat", Node="node-a", Timestamp=NOW + 100 * MS, Value=181.0)
    out, d = R.EnrichDNS(None, span, sig, R.DEFAULT_WINDOW, R.DEFAULT_ENRICHMENT_THRESHOLD)
    assert d.Matched and len(out) == 0
    out, d = R.EnrichDNS(None, R.SpanRef("trace-1", Timestamp=NOW),
                         R.SignalRef("dns_latency_ms", TraceID="trace-1", Timestamp=NOW + 200 * MS, Value=190.0),
                         R.DEFAULT_WINDOW, R.DEFAULT_ENRICHMENT_THRESHOLD)
    assert d.Matched and d.Confidence == 1.0
    assert out["llm.ebpf.dns.latency_ms"] == 190.0 and out["llm.ebpf.correlation_confidence"] == 1.0


def test_zero_time_never_matches():  # dns.go:108
    assert not R.Match(R.SpanRef("t", Timestamp=None), R.SignalRef("dns_latency_ms", TraceID="t", Timestamp=NOW), 0).Matched
    assert R.Match(R.SpanRef("t", Timestamp=0), R.SignalRef("dns_latency_ms", TraceID="t", Timestamp=0), 0).Matched


# ---- evaluator_test.go:10-73 TestEvaluateLabeledPairs
def test_evaluate_labeled_pairs():
    pairs = [
        R.LabeledPair("tp-trace", R.SpanRef("trace-1", Timestamp=NOW),
                      R.SignalRef("dns_latency_ms", TraceID="trace-1", Timestamp=NOW + 10 * MS), True, "trace_id_exact"),
        R.LabeledPair("fn-low-confidence", R.SpanRef(Service="svc-a", Node="node-a", Timestamp=NOW),
                      R.SignalRef("dns_latency_ms", Service="svc-a", Node="node-a", Timestamp=NOW + 100 * MS), True,
                      "service_node_500ms"),
        R.LabeledPair("tn-unmatched", R.SpanRef(Service="svc-a", Node="node-a", Timestamp=NOW),
                      R.SignalRef("dns_latency_ms", Service="svc-b", Node="node-b", Timestamp=NOW + 10 * MS), False),
    ]
    report, preds = R.EvaluateLabeledPairs(pairs, 2 * S, 0.7)
    assert len(preds) == 3
    assert (report.TruePositive, report.FalseNegative, report.TrueNegative) == (1, 1, 1)
    assert report.Precision == 1.0 and report.Recall == 0.5


def test_evaluate_gate():  # evaluator_test.go:75-86
    rep = R.EvalReport(Precision=0.91, Recall=0.86)
    assert R.EvaluateGate(rep, 0.9, 0.85)[0]
    assert not R.EvaluateGate(rep, 0.95, 0.85)[0]


# ---- the 55-pair fixture: the reference's own labels + SURVEY §4 golden values
def test_labeled_pairs_fixture():
    recs = load_golden("labeled_pairs")
    assert len(recs) == 55
    report, preds = R.EvaluateLabeledPairs(pairs_from_records(recs), 2000 * MS, 0.7)
    assert (report.TruePositive, report.FalsePositive, report.FalseNegative, report.TrueNegative) == (40, 0, 0, 15)
    assert report.Precision == 1.0 and report.Recall == 1.0 and report.F1 == 1.0 and report.TierAccuracy == 1.0
    assert report.MeanConfidence == 0.9199999999999997
    # CI gate of the reference: Makefile:143-151
    assert R.EvaluateGate(report, 0.90, 0.85)[0]
    hist = {}
    for p in preds:
        hist[p.Tier] = hist.get(p.Tier, 0) + 1
    assert hist == {"trace_id_exact": 17, "pod_pid_100ms": 14, "pod_conn_250ms": 9, "service_node_500ms": 9, "": 6}
    # every labelled tier of a predicted pair is reproduced
    for r, p in zip(recs, preds):
        assert p.Predicted == r["expected_match"], r["case_id"]
        if r["expected_match"] and r.get("expected_tier"):
            assert p.Tier == r["expected_tier"], r["case_id"]
    derived = load_golden("derived")["pairs"]
    assert [p.Tier for p in preds] == derived["tiers"]


# ---- retry_storm_test.go:8-110
def test_retry_storm_below_and_reach_threshold():
    d = R.RetryStormDetector(10 * S, 5)
    for i in range(4):
        assert not d.Record("pod-a", NOW + i * S)
    assert not d.IsStorm("pod-a", NOW + 4 * S)
    assert d.Record("pod-a", NOW + 4 * S)
    assert d.IsStorm("pod-a", NOW + 4 * S)


def test_retry_storm_window_expiry_pods_reset_unknown():
    d = R.RetryStormDetector(10 * S, 5)
    for _ in range(5):
        d.Record("pod-a", NOW)
    assert d.IsStorm("pod-a", NOW)
    assert not d.IsStorm("pod-a", NOW + 11 * S)
    assert d.Count("pod-a", NOW + 11 * S) == 0
    d = R.RetryStormDetector(10 * S, 3)
    for i in range(3):
        d.Record("pod-a", NOW + i * S)
    d.Record("pod-b", NOW)
    assert d.IsStorm("pod-a", NOW + 2 * S) and not d.IsStorm("pod-b", NOW + 2 * S)
    d.Reset()
    assert not d.IsStorm("pod-a", NOW) and d.Count("pod-a", NOW) == 0
    assert not d.IsStorm("nonexistent", NOW) and d.Count("nonexistent", NOW) == 0


# ---- ebpfcorrelator/correlator_test.go
def _sig(name, dt_ms, value, pid=101, trace="trace-9"):
    return R.SignalRef(name, trace, "rag", "node-a", "pod-a", pid, "", NOW + dt_ms * MS, value)


def test_enrich_dns_attributes():  # correlator_test.go:11-36
    out, d = R.New().EnrichDNSAttributes(None, R.SpanRef("trace-7", Timestamp=NOW),
                                         R.SignalRef("dns_latency_ms", TraceID="trace-7", Timestamp=NOW + 500 * MS, Value=165.0))
    assert d.Matched and out[R.ATTR["dns_latency_ms"]] == 165.0 and out[R.ATTR_CORRELATION_CONF] == 1.0


def test_enrich_attributes_multi_signal_fanout():  # correlator_test.go:38-112
    span = R.SpanRef("trace-9", "rag", "node-a", "pod-a", 101, "10.0.0.2:42424->10.0.0.53:443/tcp", NOW)
    res = R.New().EnrichAttributes(None, span, [_sig("dns_latency_ms", 10, 120), _sig("tcp_retransmits_total", 15, 5),
                                                _sig("runqueue_delay_ms", 20, 22), _sig("connect_latency_ms", 22, 180)])
    assert len(res.Candidates) == 3 and res.Debug.FanoutDropped == 1
    assert R.ATTR["dns_latency_ms"] in res.Attributes and R.ATTR["tcp_retransmits_total"] in res.Attributes
    assert R.ATTR["runqueue_delay_ms"] in res.Attributes and R.ATTR["connect_latency_ms"] not in res.Attributes


def test_enrich_attributes_threshold_and_debug():  # correlator_test.go:114-161
    span = R.SpanRef(Service="svc-a", Node="node-a", Pod="pod-a", Timestamp=NOW)
    res = R.New().EnrichAttributes(None, span, [
        R.SignalRef("dns_latency_ms", Service="svc-a", Node="node-a", Pod="pod-a", Timestamp=NOW + 300 * MS, Value=99),
        R.SignalRef("cpu_steal_pct", Service="svc-b", Node="node-z", Pod="pod-z", Timestamp=NOW + 20 * MS, Value=9),
        R.SignalRef("unknown_signal", TraceID="trace-a", Timestamp=NOW, Value=1)])
    assert len(res.Candidates) == 0
    assert (res.Debug.LowConfidence, res.Debug.Unmatched, res.Debug.UnsupportedType) == (1, 1, 1)


def test_decompose_retrieval():  # correlator_test.go:163-204
    attrs = {R.ATTR["dns_latency_ms"]: 12.5, R.ATTR["connect_latency_ms"]: 18.0, R.ATTR["tls_handshake_ms"]: 22.0,
             R.ATTR["cpu_steal_pct"]: 0.6}
    assert R.DecomposeRetrieval(attrs) == 12.5 + 18.0 + 22.0 == attrs[R.ATTR_RETRIEVAL_KERNEL_MS]
    assert R.DecomposeRetrieval({R.ATTR["dns_latency_ms"]: 15.0}) == 15.0
    attrs = {R.ATTR["cpu_steal_pct"]: 2.0}
    assert R.DecomposeRetrieval(attrs) == 0 and R.ATTR_RETRIEVAL_KERNEL_MS not in attrs


def test_process_batch_enriches_spans():  # correlator_test.go:206-260
    batch = [R.SpanRecord("trace-1", "", "rag", "node-a", "pod-a", 1001, "", NOW),
             R.SpanRecord("trace-2", "", "rag", "node-a", "pod-a", 1002, "", NOW)]
    signals = [R.SignalRef("dns_latency_ms", "trace-1", "rag", "node-a", "pod-a", 1001, "", NOW + 10 * MS, 88),
               R.SignalRef("unknown_signal", TraceID="trace-1", Timestamp=NOW, Value=1)]
    out = R.New().ProcessBatch(batch, signals)
    assert len(out.Spans) == 2
    assert out.Spans[0].Attributes[R.ATTR["dns_latency_ms"]] == 88
    assert out.Spans[0].Attributes[R.ATTR_RETRIEVAL_KERNEL_MS] == 88
    assert R.ATTR["dns_latency_ms"] not in out.Spans[1].Attributes
    assert out.Debug.UnsupportedType == 2


def test_process_batch_aggregates_fanout_drops():  # correlator_test.go:262-324
    span = R.SpanRecord("trace-9", "", "rag", "node-a", "pod-a", 222, "", NOW)
    signals = [_sig("dns_latency_ms", 5, 10, 222), _sig("connect_latency_ms", 6, 11, 222),
               _sig("tls_handshake_ms", 7, 12, 222), _sig("runqueue_delay_ms", 8, 13, 222)]
    assert R.New().ProcessBatch([span], signals).Debug.FanoutDropped == 1


# ---- pkg/attribution/bayesian_test.go
def test_posteriors_sum_to_one_and_top1():
    ba = R.NewBayesianAttributor()
    assert abs(sum(p.Posterior for p in ba.Attribute({"dns_latency_ms": 220, "connect_latency_ms": 130})) - 1.0) <= 1e-6
    cases = [
        ({"dns_latency_ms": 220, "connect_latency_ms": 130, "tcp_retransmits_total": 0.2, "runqueue_delay_ms": 4,
          "cpu_steal_pct": 0.6}, "network_dns"),
        ({"runqueue_delay_ms": 28, "cpu_steal_pct": 9, "cfs_throttled_ms": 170, "dns_latency_ms": 12}, "cpu_throttle"),
        ({"cfs_throttled_ms": 90, "mem_reclaim_latency_ms": 25, "disk_io_latency_ms": 60, "runqueue_delay_ms": 14,
          "dns_latency_ms": 12}, "memory_pressure"),
        ({"connect_latency_ms": 95, "tls_handshake_ms": 90, "syscall_latency_ms": 300, "connect_errors_total": 2,
          "dns_latency_ms": 12}, "provider_throttle"),
    ]
    for signals, want in cases:
        post = ba.Attribute(signals)
        assert post[0].Domain == want
    assert ba.Attribute(cases[0][0])[0].Posterior >= 0.3
    assert len(R.ALL_DOMAINS) == 8 and all(v == 1 / 8 for v in R.DefaultPriors().values())


def test_no_signals_falls_back_to_rule():  # bayesian_test.go (AttributeSample without Signals)
    ba = R.NewBayesianAttributor()
    out = ba.AttributeSample(R.FaultSample("i", "dns_latency", Confidence=0.9))
    assert out.PredictedFaultDomain == "network_dns" and out.Confidence == 0.9 and out.FaultHypotheses == []


def test_map_fault_label():  # mapper_test.go
    assert R.MapFaultLabel("network_partition") == "network_egress"
    assert R.MapFaultLabel("dns_latency") == "network_dns"
    assert R.MapFaultLabel("something_else") == "unknown"


# ---- pkg/attribution/pipeline_test.go:46-175
def _pred(dom, hyps=()):
    return R.IncidentAttribution("", dom, 0.0, [(d, p, []) for d, p in hyps])


def test_confusion_and_accuracies():
    m = R.BuildConfusionMatrix([R.FaultSample("inc-1", "provider_throttle", "network_dns")], [_pred("provider_throttle")])
    assert m[("network_dns", "provider_throttle")] == 1
    assert R.Accuracy([R.FaultSample(FaultLabel="provider_throttle"), R.FaultSample(FaultLabel="dns_latency")],
                      [_pred("provider_throttle"), _pred("provider_throttle")]) == 0.5
    assert R.PartialAccuracy([R.FaultSample(FaultLabel="dns_latency", ExpectedDomains=["network_dns"]),
                              R.FaultSample(FaultLabel="cpu_throttle", ExpectedDomains=["cpu_throttle"])],
                             [_pred("network_dns"), _pred("cpu_throttle")]) == 1.0
    assert R.PartialAccuracy([R.FaultSample(ExpectedDomains=["network_dns", "cpu_throttle"]),
                              R.FaultSample(ExpectedDomains=["provider_throttle", "memory_pressure"])],
                             [_pred("cpu_throttle"), _pred("network_dns")]) == 0.5
    s = [R.FaultSample(ExpectedDomains=["network_dns", "cpu_throttle"])]
    assert R.CoverageAccuracy(s, [_pred("network_dns", [("network_dns", 0.5), ("cpu_throttle", 0.3)])], 0.1) == 1.0
    assert R.CoverageAccuracy(s, [_pred("network_dns", [("network_dns", 0.5), ("memory_pressure", 0.3)])], 0.1) == 0.5
    assert R.CoverageAccuracy(s, [_pred("network_dns", [("network_dns", 0.5), ("cpu_throttle", 0.02)])], 0.05) == 0.5


# ---- the 55-incident multi-fault fixture vs SURVEY §4 golden values
def test_multi_fault_fixture():
    samples = samples_from_records(load_golden("multi_fault_samples"))
    assert len(samples) == 55
    attr = R.BuildAttributionsBayes(samples, None)
    assert R.Accuracy(samples, attr) == 29 / 55
    assert R.PartialAccuracy(samples, attr) == 54 / 55
    assert abs(R.CoverageAccuracy(samples, attr, 0.10) - 0.8484848484848485) < 1e-12
    want = {("cpu_throttle", "cpu_throttle"): 6, ("memory_pressure", "memory_pressure"): 6,
            ("network_dns", "network_dns"): 7, ("network_egress", "network_egress"): 5,
            ("provider_throttle", "provider_error"): 1, ("provider_throttle", "provider_throttle"): 5,
            ("unknown", "cpu_throttle"): 8, ("unknown", "memory_pressure"): 11, ("unknown", "network_egress"): 3,
            ("unknown", "provider_throttle"): 3}
    assert R.BuildConfusionMatrix(samples, attr) == want
    hyp = [dict((d, p) for d, p, _ in a.FaultHypotheses) for a in attr]
    assert abs(hyp[0]["network_dns"] - 0.933270) < 5e-7           # mf-01
    assert abs(attr[1].Confidence - 0.996371) < 5e-7               # mf-02 cpu_throttle
    assert abs(hyp[29]["provider_throttle"] - 0.326321) < 5e-7     # mf-30
    assert abs(hyp[29]["retrieval_backend"] - 0.306570) < 5e-7
    for a in attr:
        assert abs(sum(p for _, p, _ in a.FaultHypotheses) - 1.0) < 0.08  # only >= 0.01 hypotheses are listed
    derived = load_golden("derived")["multi_fault"]
    assert [a.PredictedFaultDomain for a in attr] == derived["predicted"]


def test_mixed_fault_fixture_rule_path():  # pkg/benchmark/harness_test.go:55-66 input: label-only samples
    samples = samples_from_records(load_golden("mixed_fault_samples"))
    attr = R.BuildAttributionsBayes(samples, None)
    assert all(a.FaultHypotheses == [] for a in attr)
    assert R.Accuracy(samples, attr) == 1.0
