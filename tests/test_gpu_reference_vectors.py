"""GPU parity, part 1: the reference's own unit tests and fixtures, run through the host mirror of the
Go interfaces (same names and arguments) over the C ABI.  Transcribed from pkg/correlation/dns_test.go,
evaluator_test.go, ebpfcorrelator/correlator_test.go, pkg/attribution/bayesian_test.go, pipeline_test.go.
Bit-exact for tiers, match sets, counters and labels; posteriors within 1e-6 relative (north_star)."""
import math

import pytest

from llm_slo_ebpf_toolkit_b200 import attribution as AT
from llm_slo_ebpf_toolkit_b200 import correlation as CO
from llm_slo_ebpf_toolkit_b200 import ebpfcorrelator as EC
from oracle import ref_str as R
from tests.fixtures import load_golden, parse_ts

pytestmark = pytest.mark.gpu
NOW = 1_771_372_800_123_456_789
MS, S = CO.MS, CO.SECOND


def test_match_tiers(ctx):  # dns_test.go:8-87
    span = CO.SpanRef("trace-1", "chat", "node-a", "pod-a", 42, "10.0.0.1:1234->10.0.0.2:53/udp", NOW)
    cases = [
        (CO.SignalRef("dns_latency_ms", TraceID="trace-1", Timestamp=NOW + 1 * S), 1.0, "trace_id_exact"),
        (CO.SignalRef("dns_latency_ms", Pod="pod-a", PID=42, Timestamp=NOW + 50 * MS), 0.9, "pod_pid_100ms"),
        (CO.SignalRef("dns_latency_ms", Pod="pod-a", ConnTuple="10.0.0.1:1234->10.0.0.2:53/udp", Timestamp=NOW + 200 * MS),
         0.8, "pod_conn_250ms"),
        (CO.SignalRef("dns_latency_ms", Service="chat", Node="node-a", Timestamp=NOW + 400 * MS), 0.65, "service_node_500ms"),
    ]
    for sig, conf, tier in cases:
        d = CO.Match(span, sig, CO.DefaultWindow, ctx)
        assert d.Matched and d.Confidence == conf and d.Tier == tier


def test_enrich_dns_respects_threshold(ctx):  # dns_test.go:89-132
    out, d = CO.EnrichDNS(None, CO.SpanRef(Service="chat", Node="node-a", Timestamp=NOW),
                          CO.SignalRef("dns_latency_ms", Service="chat", Node="node-a", Timestamp=NOW + 100 * MS, Value=181.0),
                          CO.DefaultWindow, CO.DefaultEnrichmentThreshold, ctx)
    assert d.Matched and len(out) == 0
    out, d = CO.EnrichDNS(None, CO.SpanRef("trace-1", Timestamp=NOW),
                          CO.SignalRef("dns_latency_ms", TraceID="trace-1", Timestamp=NOW + 200 * MS, Value=190.0),
                          CO.DefaultWindow, CO.DefaultEnrichmentThreshold, ctx)
    assert d.Matched and d.Confidence == 1.0
    assert out["llm.ebpf.dns.latency_ms"] == 190.0 and out["llm.ebpf.correlation_confidence"] == 1.0


def test_zero_time_and_unix_zero(ctx):  # dns.go:108: IsZero is year 1, not Unix 0
    assert not CO.Match(CO.SpanRef("t", Timestamp=None), CO.SignalRef("dns_latency_ms", TraceID="t", Timestamp=NOW), 0, ctx).Matched
    assert CO.Match(CO.SpanRef("t", Timestamp=0), CO.SignalRef("dns_latency_ms", TraceID="t", Timestamp=0), 0, ctx).Matched


def test_evaluate_labeled_pairs(ctx):  # evaluator_test.go:10-73
    pairs = [
        CO.LabeledPair("tp-trace", CO.SpanRef("trace-1", Timestamp=NOW),
                       CO.SignalRef("dns_latency_ms", TraceID="trace-1", Timestamp=NOW + 10 * MS), True, "trace_id_exact"),
        CO.LabeledPair("fn-low-confidence", CO.SpanRef(Service="svc-a", Node="node-a", Timestamp=NOW),
                       CO.SignalRef("dns_latency_ms", Service="svc-a", Node="node-a", Timestamp=NOW + 100 * MS), True,
                       "service_node_500ms"),
        CO.LabeledPair("tn-unmatched", CO.SpanRef(Service="svc-a", Node="node-a", Timestamp=NOW),
                       CO.SignalRef("dns_latency_ms", Service="svc-b", Node="node-b", Timestamp=NOW + 10 * MS), False),
    ]
    report, preds = CO.EvaluateLabeledPairs(pairs, 2 * S, 0.7, ctx)
    assert len(preds) == 3
    assert (report.TruePositive, report.FalseNegative, report.TrueNegative) == (1, 1, 1)
    assert report.Precision == 1.0 and report.Recall == 0.5
    assert CO.EvaluateGate(CO.EvalReport(Precision=0.91, Recall=0.86), 0.9, 0.85).Pass   # evaluator_test.go:75-86
    assert not CO.EvaluateGate(CO.EvalReport(Precision=0.91, Recall=0.86), 0.95, 0.85).Pass
    assert CO.EvaluateLabeledPairs([], 0, 0, ctx)[1] == []


def _pairs():
    out = []
    for r in load_golden("labeled_pairs"):
        sp, sg = r["span"], r["signal"]
        out.append(CO.LabeledPair(
            r["case_id"],
            CO.SpanRef(sp.get("trace_id", ""), sp.get("service", ""), sp.get("node", ""), sp.get("pod", ""),
                       sp.get("pid", 0), sp.get("conn_tuple", ""), parse_ts(sp.get("timestamp"))),
            CO.SignalRef(sg["signal"], sg.get("trace_id", ""), sg.get("service", ""), sg.get("node", ""), sg.get("pod", ""),
                         sg.get("pid", 0), sg.get("conn_tuple", ""), parse_ts(sg.get("timestamp")), sg.get("value", 0.0)),
            r["expected_match"], r.get("expected_tier", "")))
    return out


def test_labeled_pairs_fixture_ci_gate(ctx):  # Makefile:143-151 on pkg/correlation/testdata/labeled_pairs.jsonl
    pairs = _pairs()
    report, preds = CO.EvaluateLabeledPairs(pairs, 2000 * MS, 0.7, ctx)
    assert (report.TruePositive, report.FalsePositive, report.FalseNegative, report.TrueNegative) == (40, 0, 0, 15)
    assert report.Precision == 1.0 and report.Recall == 1.0 and report.F1 == 1.0 and report.TierAccuracy == 1.0
    assert abs(report.MeanConfidence - 0.92) <= 1e-6 * 0.92
    assert CO.EvaluateGate(report, 0.90, 0.85).Pass
    derived = load_golden("derived")["pairs"]
    assert [p.Tier for p in preds] == derived["tiers"]
    assert [p.Predicted for p in preds] == derived["predicted"]
    # sweep of windows and thresholds (config 5) against the string oracle
    from tests.fixtures import pairs_from_records
    rp = pairs_from_records(load_golden("labeled_pairs"))
    for window_ms in (50, 100, 250, 500, 1000, 2000, 5000):
        for thr in (0.5, 0.65, 0.7, 0.8, 0.9, 1.0):
            got, gp = CO.EvaluateLabeledPairs(pairs, window_ms * MS, thr, ctx)
            want, wp = R.EvaluateLabeledPairs(rp, window_ms * MS, thr)
            assert [p.Tier for p in gp] == [p.Tier for p in wp]
            assert (got.TruePositive, got.FalsePositive, got.FalseNegative, got.TrueNegative) == \
                (want.TruePositive, want.FalsePositive, want.FalseNegative, want.TrueNegative)
            assert got.TierAccuracy == want.TierAccuracy and got.Precision == want.Precision and got.Recall == want.Recall
            assert abs(got.MeanConfidence - want.MeanConfidence) <= 1e-6 * max(want.MeanConfidence, 1e-300)


def _sig(name, dt_ms, value, pid=101, trace="trace-9"):
    return CO.SignalRef(name, trace, "rag", "node-a", "pod-a", pid, "", NOW + dt_ms * MS, value)


def test_enrich_attributes_multi_signal_fanout(ctx):  # correlator_test.go:38-112
    span = CO.SpanRef("trace-9", "rag", "node-a", "pod-a", 101, "10.0.0.2:42424->10.0.0.53:443/tcp", NOW)
    res = EC.New(ctx).EnrichAttributes(None, span, [_sig("dns_latency_ms", 10, 120), _sig("tcp_retransmits_total", 15, 5),
                                                    _sig("runqueue_delay_ms", 20, 22), _sig("connect_latency_ms", 22, 180)])
    assert len(res.Candidates) == 3 and res.Debug.FanoutDropped == 1
    assert "llm.ebpf.dns.latency_ms" in res.Attributes and "llm.ebpf.tcp.retransmits" in res.Attributes
    assert "llm.ebpf.sched.runqueue_delay_ms" in res.Attributes
    assert "llm.ebpf.net.connect_latency_ms" not in res.Attributes   # the farthest in time is dropped
    assert [c.Index for c in res.Candidates] == [0, 1, 2]


def test_enrich_attributes_threshold_and_debug(ctx):  # correlator_test.go:114-161
    span = CO.SpanRef(Service="svc-a", Node="node-a", Pod="pod-a", Timestamp=NOW)
    res = EC.New(ctx).EnrichAttributes(None, span, [
        CO.SignalRef("dns_latency_ms", Service="svc-a", Node="node-a", Pod="pod-a", Timestamp=NOW + 300 * MS, Value=99),
        CO.SignalRef("cpu_steal_pct", Service="svc-b", Node="node-z", Pod="pod-z", Timestamp=NOW + 20 * MS, Value=9),
        CO.SignalRef("unknown_signal", TraceID="trace-a", Timestamp=NOW, Value=1)])
    assert len(res.Candidates) == 0
    assert (res.Debug.LowConfidence, res.Debug.Unmatched, res.Debug.UnsupportedType) == (1, 1, 1)


def test_enrich_dns_attributes(ctx):  # correlator_test.go:11-36
    out, d = EC.New(ctx).EnrichDNSAttributes(None, CO.SpanRef("trace-7", Timestamp=NOW),
                                             CO.SignalRef("dns_latency_ms", TraceID="trace-7", Timestamp=NOW + 500 * MS, Value=165.0))
    assert d.Matched and out["llm.ebpf.dns.latency_ms"] == 165.0 and out["llm.ebpf.correlation_confidence"] == 1.0


def test_process_batch_enriches_spans(ctx):  # correlator_test.go:206-260
    batch = [EC.SpanRecord("trace-1", "", "rag", "node-a", "pod-a", 1001, "", NOW),
             EC.SpanRecord("trace-2", "", "rag", "node-a", "pod-a", 1002, "", NOW)]
    signals = [CO.SignalRef("dns_latency_ms", "trace-1", "rag", "node-a", "pod-a", 1001, "", NOW + 10 * MS, 88),
               CO.SignalRef("unknown_signal", TraceID="trace-1", Timestamp=NOW, Value=1)]
    out = EC.New(ctx).ProcessBatch(batch, signals)
    assert len(out.Spans) == 2
    assert out.Spans[0].Attributes["llm.ebpf.dns.latency_ms"] == 88
    assert out.Spans[0].Attributes["llm.ebpf.retrieval.kernel_attributed_ms"] == 88
    assert "llm.ebpf.dns.latency_ms" not in out.Spans[1].Attributes
    assert out.Debug.UnsupportedType == 2
    assert EC.New(ctx).ProcessBatch([], signals).Spans == []


def test_process_batch_aggregates_fanout_drops(ctx):  # correlator_test.go:262-324
    span = EC.SpanRecord("trace-9", "", "rag", "node-a", "pod-a", 222, "", NOW)
    signals = [_sig("dns_latency_ms", 5, 10, 222), _sig("connect_latency_ms", 6, 11, 222),
               _sig("tls_handshake_ms", 7, 12, 222), _sig("runqueue_delay_ms", 8, 13, 222)]
    out = EC.New(ctx).ProcessBatch([span], signals)
    assert out.Debug.FanoutDropped == 1
    # DecomposeRetrieval over the kept three: dns + connect + tls (correlator_test.go:163-204)
    assert out.Spans[0].Attributes["llm.ebpf.retrieval.kernel_attributed_ms"] == 10 + 11 + 12


def test_bayesian_unit_vectors(ctx):  # bayesian_test.go:9-195
    ba = AT.NewBayesianAttributor(ctx)
    post = ba.Attribute({"dns_latency_ms": 220, "connect_latency_ms": 130})
    assert abs(sum(p.Posterior for p in post) - 1.0) <= 1e-6 and len(post) == 8
    cases = [
        ({"dns_latency_ms": 220, "connect_latency_ms": 130, "tcp_retransmits_total": 0.2, "runqueue_delay_ms": 4,
          "cpu_steal_pct": 0.6}, "network_dns"),
        ({"runqueue_delay_ms": 28, "cpu_steal_pct": 9, "cfs_throttled_ms": 170, "dns_latency_ms": 12}, "cpu_throttle"),
        ({"cfs_throttled_ms": 90, "mem_reclaim_latency_ms": 25, "disk_io_latency_ms": 60, "runqueue_delay_ms": 14,
          "dns_latency_ms": 12}, "memory_pressure"),
        ({"connect_latency_ms": 95, "tls_handshake_ms": 90, "syscall_latency_ms": 300, "connect_errors_total": 2,
          "dns_latency_ms": 12}, "provider_throttle"),
    ]
    rb = R.NewBayesianAttributor()
    for signals, want in cases:
        got = ba.Attribute(signals)
        ref = rb.Attribute(signals)
        assert got[0].Domain == want
        assert [p.Domain for p in got] == [p.Domain for p in ref]
        for g, w in zip(got, ref):
            assert abs(g.Posterior - w.Posterior) <= 1e-6 * w.Posterior
            assert g.Evidence == w.Evidence
    assert ba.Attribute(cases[0][0])[0].Posterior >= 0.3
    assert all(v == 1 / 8 for v in AT.DefaultPriors().values()) and len(AT.AllDomains()) == 8
    out = ba.AttributeSample(AT.FaultSample("i", FaultLabel="dns_latency", Confidence=0.9))   # no Signals -> rule path
    assert out.PredictedFaultDomain == "network_dns" and out.Confidence == 0.9 and out.FaultHypotheses == []
    # a mutated model takes effect (Priors / Likelihoods are public fields in Go)
    ba.Priors["cpu_throttle"] = 0.0
    del ba.Likelihoods["dns_latency_ms"]
    ba.Likelihoods["custom_signal"] = {"network_dns": 0.3}
    rb.Priors["cpu_throttle"] = 0.0
    del rb.Likelihoods["dns_latency_ms"]
    rb.Likelihoods["custom_signal"] = {"network_dns": 0.3}
    for signals, _ in cases:
        for g, w in zip(sorted(ba.Attribute(signals), key=lambda p: p.Domain), sorted(rb.Attribute(signals), key=lambda p: p.Domain)):
            assert abs(g.Posterior - w.Posterior) <= 1e-6 * w.Posterior + 1e-300


def _pred(dom, hyps=()):
    return AT.IncidentAttribution(PredictedFaultDomain=dom, FaultHypotheses=[AT.FaultHypothesis(d, p, []) for d, p in hyps])


def test_pipeline_unit_vectors(ctx):  # pipeline_test.go:46-175
    m = AT.BuildConfusionMatrix([AT.FaultSample("inc-1", FaultLabel="provider_throttle", ExpectedDomain="network_dns")],
                                [_pred("provider_throttle")], ctx)
    assert m == {AT.MatrixKey("network_dns", "provider_throttle"): 1}
    assert AT.Accuracy([AT.FaultSample(FaultLabel="provider_throttle"), AT.FaultSample(FaultLabel="dns_latency")],
                       [_pred("provider_throttle"), _pred("provider_throttle")], ctx) == 0.5
    assert AT.PartialAccuracy([AT.FaultSample(FaultLabel="dns_latency", ExpectedDomains=["network_dns"]),
                               AT.FaultSample(FaultLabel="cpu_throttle", ExpectedDomains=["cpu_throttle"])],
                              [_pred("network_dns"), _pred("cpu_throttle")], ctx) == 1.0
    assert AT.PartialAccuracy([AT.FaultSample(ExpectedDomains=["network_dns", "cpu_throttle"]),
                               AT.FaultSample(ExpectedDomains=["provider_throttle", "memory_pressure"])],
                              [_pred("cpu_throttle"), _pred("network_dns")], ctx) == 0.5
    s = [AT.FaultSample(ExpectedDomains=["network_dns", "cpu_throttle"])]
    assert AT.CoverageAccuracy(s, [_pred("network_dns", [("network_dns", 0.5), ("cpu_throttle", 0.3)])], 0.1, ctx) == 1.0
    assert AT.CoverageAccuracy(s, [_pred("network_dns", [("network_dns", 0.5), ("memory_pressure", 0.3)])], 0.1, ctx) == 0.5
    assert AT.CoverageAccuracy(s, [_pred("network_dns", [("network_dns", 0.5), ("cpu_throttle", 0.02)])], 0.05, ctx) == 0.5
    assert AT.Accuracy([], [], ctx) == 0 and AT.BuildConfusionMatrix([], [], ctx) == {}
    # actual strings outside AllDomains get their own rows
    m = AT.BuildConfusionMatrix([AT.FaultSample(ExpectedDomain="weird"), AT.FaultSample(ExpectedDomain="weird")],
                                [_pred("unknown"), _pred("network_dns")], ctx)
    assert m == {AT.MatrixKey("weird", "unknown"): 1, AT.MatrixKey("weird", "network_dns"): 1}


def _samples(name):
    return [AT.FaultSample(r.get("incident_id", ""), r.get("timestamp"), r.get("cluster", ""), r.get("namespace", ""),
                           r.get("service", ""), r.get("fault_label", ""), r.get("expected_domain", ""),
                           list(r.get("expected_domains") or []), dict(r.get("signals") or {}), r.get("confidence", 0.0),
                           r.get("burn_rate", 0.0), r.get("window_minutes", 0), r.get("request_id", ""), r.get("trace_id", ""))
            for r in load_golden(name)]


def test_multi_fault_fixture(ctx):  # pkg/attribution/testdata/multi_fault_samples.jsonl, golden values of SURVEY.md §4
    samples = _samples("multi_fault_samples")
    attr = AT.BuildAttributions(samples, "bayes")
    derived = load_golden("derived")["multi_fault"]
    assert [a.PredictedFaultDomain for a in attr] == derived["predicted"]          # argmax labels: bit-exact
    for a, conf, hyp in zip(attr, derived["confidence"], derived["hypotheses"]):
        assert abs(a.Confidence - conf) <= 1e-6 * conf                               # posteriors: 1e-6 relative
        assert {h.Domain for h in a.FaultHypotheses} == set(hyp)
        for h in a.FaultHypotheses:
            assert abs(h.Posterior - hyp[h.Domain]) <= 1e-6 * hyp[h.Domain]
    assert AT.Accuracy(samples, attr, ctx) == 29 / 55
    assert AT.PartialAccuracy(samples, attr, ctx) == 54 / 55
    assert abs(AT.CoverageAccuracy(samples, attr, 0.10, ctx) - derived["coverage_0.10"]) <= 1e-6
    got = sorted([k.Actual, k.Predicted, v] for k, v in AT.BuildConfusionMatrix(samples, attr, ctx).items())
    assert got == derived["confusion"]
    assert abs(attr[0].FaultHypotheses[0].Posterior - 0.933270) < 5e-7               # mf-01
    assert attr[0].FaultHypotheses[0].Evidence == ["connect_latency_ms", "dns_latency_ms"]
    # rule mode and the label-only fixture (pkg/benchmark/testdata/mixed_fault_samples.jsonl)
    mixed = _samples("mixed_fault_samples")
    assert AT.Accuracy(mixed, AT.BuildAttributions(mixed, "bayes"), ctx) == 1.0
    assert [a.PredictedFaultDomain for a in AT.BuildAttributions(samples, " RULE ")] == \
        [AT.MapFaultLabel(s.FaultLabel) for s in samples]
