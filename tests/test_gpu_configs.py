"""BASELINE.json configs[0] and configs[4] as parity cases through the host mirror.

C1: cmd/faultreplay's single-fault scenario (DNS latency) through correlation + attribution.  The reference has no
    glue between the two: samples from pkg/faultreplay/generator.go:20-58, the twelve signal values of
    pkg/signals/generator.go:244-289 (profileForFault) as events at the sample's timestamp and trace id, one span
    per sample; ProcessBatch enriches the span, the enriched attributes become FaultSample.Signals, Attribute labels
    the incident.  Expected: network_dns at posterior 0.933270.
C5: the fault-matrix sweep: every replay scenario x injection intensity, batch-attributed, against the string oracle.
"""
import math

import pytest

from llm_slo_ebpf_toolkit_b200 import attribution as AT
from llm_slo_ebpf_toolkit_b200 import correlation as CO
from llm_slo_ebpf_toolkit_b200 import ebpfcorrelator as EC
from oracle import ref_str as R

pytestmark = pytest.mark.gpu
START = 1_771_372_800_000_000_000
S = CO.SECOND

BASE = {"dns_latency_ms": 12, "tcp_retransmits_total": 0.2, "runqueue_delay_ms": 4, "connect_latency_ms": 18,
        "connect_errors_total": 0, "tls_handshake_ms": 22, "tls_handshake_fail_total": 0, "cpu_steal_pct": 0.6,
        "cfs_throttled_ms": 5, "mem_reclaim_latency_ms": 0.5, "disk_io_latency_ms": 2, "syscall_latency_ms": 5}
FAULTS = {   # profileForFault, pkg/signals/generator.go:262-287
    "dns_latency": {"dns_latency_ms": 220, "connect_latency_ms": 130},
    "cpu_throttle": {"runqueue_delay_ms": 28, "cpu_steal_pct": 9, "cfs_throttled_ms": 170},
    "memory_pressure": {"runqueue_delay_ms": 14, "cfs_throttled_ms": 90, "mem_reclaim_latency_ms": 25, "disk_io_latency_ms": 60},
    "provider_throttle": {"connect_latency_ms": 45, "tls_handshake_ms": 55, "connect_errors_total": 1, "syscall_latency_ms": 250},
    "network_partition": {"connect_latency_ms": 350, "connect_errors_total": 3, "tcp_retransmits_total": 12, "dns_latency_ms": 180,
                          "tls_handshake_fail_total": 2},
}
SCENARIOS = {   # pkg/faultreplay/generator.go:10-17
    "provider_throttle": ["provider_throttle"], "dns_latency": ["dns_latency"], "cpu_throttle": ["cpu_throttle"],
    "memory_pressure": ["memory_pressure"], "network_partition": ["network_partition"],
    "mixed": ["provider_throttle", "dns_latency", "cpu_throttle", "memory_pressure", "network_partition"],
}


def profile(label, intensity=1.0):
    p = dict(BASE)
    for k, v in FAULTS.get(label, {}).items():
        p[k] = BASE[k] + (v - BASE[k]) * intensity          # injection intensity scales the excursion from the baseline
    return p


def replay(scenario, count):
    """GenerateFaultSamples (generator.go:20-58): label idx % len, timestamp start + idx s, ids replay-*-%04d."""
    labels = SCENARIOS[scenario]
    return [AT.FaultSample("replay-inc-%04d" % (i + 1), None, "local", "default", "chat", labels[i % len(labels)],
                           AT.MapFaultLabel(labels[i % len(labels)]), [], {}, 0.9, 2.0, 5, "replay-req-%04d" % (i + 1),
                           "replay-trace-%04d" % (i + 1)) for i in range(count)]


def through_correlation(samples, ctx, intensity=1.0):
    """events + spans from the samples -> ProcessBatch -> enriched attributes -> FaultSample.Signals"""
    spans, signals = [], []
    for i, s in enumerate(samples):
        ts = START + i * S
        spans.append(EC.SpanRecord(s.TraceID, "", "chat", "node-a", "pod-a", 4242, "10.244.0.10:42424->10.244.0.53:443/tcp", ts))
        for name, v in profile(s.FaultLabel, intensity).items():
            signals.append(CO.SignalRef(name, s.TraceID, "chat", "node-a", "pod-a", 4242, "", ts, float(v)))
    corr = EC.Correlator(2000, 0.7, 12, ctx)        # keep all twelve signals of a span (the default fanout of 3 keeps three)
    batch = corr.ProcessBatch(spans, signals)
    key_to_sig = {k: n for n, k in zip(__import__("llm_slo_ebpf_toolkit_b200")._abi.SIGNALS, EC.ATTR_KEYS)}
    out = []
    for s, sp in zip(samples, batch.Spans):
        sig = {key_to_sig[k]: v for k, v in sp.Attributes.items() if k in key_to_sig}
        out.append(AT.FaultSample(s.IncidentID, None, s.Cluster, s.Namespace, s.Service, s.FaultLabel, s.ExpectedDomain, [], sig,
                                  s.Confidence, s.BurnRate, s.WindowMinutes, s.RequestID, s.TraceID))
    return out, batch


def test_c1_faultreplay_dns_latency(ctx):
    samples = replay("dns_latency", 30)
    enriched, batch = through_correlation(samples, ctx)
    # one span per sample, its own twelve events and nobody else's (trace ids differ, neighbours are 1 s away: pod tiers
    # see them only within 100/250 ms)
    assert batch.Debug.FanoutDropped == 0 and all(len(e.Signals) == 12 for e in enriched)
    assert all(sp.Attributes["llm.ebpf.correlation_confidence"] == 1.0 for sp in batch.Spans)
    assert batch.Spans[0].Attributes["llm.ebpf.retrieval.kernel_attributed_ms"] == 220 + 130 + 22
    attr = AT.BuildAttributionsBayes(enriched, AT.NewBayesianAttributor(ctx))
    for a in attr:
        assert a.PredictedFaultDomain == "network_dns"
        assert abs(a.Confidence - 0.933270) < 5e-7
    assert AT.Accuracy(enriched, attr, ctx) == 1.0
    assert AT.BuildConfusionMatrix(enriched, attr, ctx) == {AT.MatrixKey("network_dns", "network_dns"): 30}
    # the default fanout keeps three signals per span: dns, tcp and runqueue are the first three in input order at |dt| = 0
    e3, b3 = through_correlation(samples[:4], ctx)
    corr3 = EC.New(ctx).ProcessBatch([EC.SpanRecord(s.TraceID, "", "chat", "node-a", "pod-a", 4242, "", START + i * S)
                                      for i, s in enumerate(samples[:4])],
                                     [CO.SignalRef(n, s.TraceID, "chat", "node-a", "pod-a", 4242, "", START + i * S, float(v))
                                      for i, s in enumerate(samples[:4]) for n, v in profile("dns_latency").items()])
    assert corr3.Debug.FanoutDropped == 4 * 9
    assert sorted(corr3.Spans[0].Attributes) == sorted(["llm.ebpf.dns.latency_ms", "llm.ebpf.tcp.retransmits",
                                                        "llm.ebpf.sched.runqueue_delay_ms", "llm.ebpf.correlation_confidence",
                                                        "llm.ebpf.retrieval.kernel_attributed_ms"])


def test_c5_fault_matrix_sweep(ctx):
    ba, rb = AT.NewBayesianAttributor(ctx), R.NewBayesianAttributor()
    samples, ref_samples = [], []
    for scenario in SCENARIOS:
        for intensity in (0.25, 0.5, 1.0, 2.0, 4.0):
            for s in replay(scenario, 10):
                sig = {k: float(v) for k, v in profile(s.FaultLabel, intensity).items()}
                samples.append(AT.FaultSample(s.IncidentID, None, "", "", "", s.FaultLabel, s.ExpectedDomain, [], sig, 0.9))
                ref_samples.append(R.FaultSample(s.IncidentID, s.FaultLabel, s.ExpectedDomain, [], sig, 0.9))
    got = AT.BuildAttributionsBayes(samples, ba)          # one launch for the whole matrix
    want = R.BuildAttributionsBayes(ref_samples, rb)
    assert len(got) == 6 * 5 * 10
    for g, w in zip(got, want):
        assert g.PredictedFaultDomain == w.PredictedFaultDomain
        assert abs(g.Confidence - w.Confidence) <= 1e-6 * w.Confidence
        assert [h.Domain for h in g.FaultHypotheses] == [d for d, _, _ in w.FaultHypotheses]
    assert AT.Accuracy(samples, got, ctx) == R.Accuracy(ref_samples, want)
    gm = {(k.Actual, k.Predicted): v for k, v in AT.BuildConfusionMatrix(samples, got, ctx).items()}
    assert gm == R.BuildConfusionMatrix(ref_samples, want)
    full = [g for g, s in zip(got, samples) if True]
    assert sum(1 for g, s in zip(got, samples) if g.PredictedFaultDomain == s.ExpectedDomain) >= 150   # most of the matrix is attributed
