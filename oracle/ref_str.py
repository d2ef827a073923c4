"""String-keyed CPU restatement of the reference's correlation + attribution path.

TEST INFRASTRUCTURE ONLY.  This module is the checker: only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import it.  Nothing under
llm-slo-ebpf-toolkit_b200/ imports it and the product path has no CPU fallback.

It restates the Go reference function by function on the reference's own string-keyed types, so it
can be pinned directly by the reference's fixtures and unit-test vectors (tests/golden/).  Pure
Python loops: small cases only.  The packed-record C++ oracle (oracle/oracle.cpp) is proven equal to
this one through the interner (tests/test_oracle_packed.py) and is what the GPU is compared against
at size.  Parity status: PINNED by pkg/correlation/testdata/labeled_pairs.jsonl (55 labelled pairs),
pkg/attribution/testdata/multi_fault_samples.jsonl and the assertions of the reference's *_test.go
files transcribed in tests/test_oracle_str.py (CPU) and tests/test_gpu_reference_vectors.py (the same vectors through the GPU).

Every function cites the reference file:line it follows (paths relative to the reference root).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

# ---------------------------------------------------------------------------------------------
# time: Go time.Time is modelled as Unix nanoseconds (int) with None for the zero Time.
# ---------------------------------------------------------------------------------------------
MS = 1_000_000
SECOND = 1_000_000_000
_DUR_MAX = (1 << 63) - 1
_DUR_MIN = -(1 << 63)


def _sub(a: int, b: int) -> int:
    """time.Time.Sub saturates to the Duration range."""
    d = a - b
    return max(_DUR_MIN, min(_DUR_MAX, d))


# pkg/correlation/dns.go:10-16
DEFAULT_WINDOW = 2 * SECOND
DEFAULT_ENRICHMENT_THRESHOLD = 0.7

# pkg/semconv/llm_ebpf.go:3-27
ATTR = {
    "dns_latency_ms": "llm.ebpf.dns.latency_ms",
    "tcp_retransmits_total": "llm.ebpf.tcp.retransmits",
    "runqueue_delay_ms": "llm.ebpf.sched.runqueue_delay_ms",
    "connect_latency_ms": "llm.ebpf.net.connect_latency_ms",
    "tls_handshake_ms": "llm.ebpf.tls.handshake_ms",
    "cpu_steal_pct": "llm.ebpf.cpu.steal_pct",
    "cfs_throttled_ms": "llm.ebpf.cpu.cfs_throttled_ms",
    "mem_reclaim_latency_ms": "llm.ebpf.mm.reclaim_latency_ms",
    "disk_io_latency_ms": "llm.ebpf.blk.io_latency_ms",
    "syscall_latency_ms": "llm.ebpf.syscall.latency_ms",
    "connect_errors_total": "llm.ebpf.net.connect_errors_total",
    "tls_handshake_fail_total": "llm.ebpf.tls.handshake_fail_total",
}
ATTR_CORRELATION_CONF = "llm.ebpf.correlation_confidence"
ATTR_RETRIEVAL_KERNEL_MS = "llm.ebpf.retrieval.kernel_attributed_ms"


@dataclass
class SpanRef:  # dns.go:19-27
    TraceID: str = ""
    Service: str = ""
    Node: str = ""
    Pod: str = ""
    PID: int = 0
    ConnTuple: str = ""
    Timestamp: Optional[int] = None


@dataclass
class SignalRef:  # dns.go:30-40
    Signal: str = ""
    TraceID: str = ""
    Service: str = ""
    Node: str = ""
    Pod: str = ""
    PID: int = 0
    ConnTuple: str = ""
    Timestamp: Optional[int] = None
    Value: float = 0.0


@dataclass
class Decision:  # dns.go:43-47
    Matched: bool = False
    Confidence: float = 0.0
    Tier: str = ""


def within_window(a: Optional[int], b: Optional[int], window: int) -> bool:
    """dns.go:107-113"""
    if a is None or b is None:
        return False
    diff = abs(float(_sub(a, b)))
    return diff <= float(window)


def Match(span: SpanRef, signal: SignalRef, window: int) -> Decision:
    """dns.go:50-76"""
    if window <= 0:
        window = DEFAULT_WINDOW
    if not within_window(span.Timestamp, signal.Timestamp, window):
        return Decision()
    if span.TraceID != "" and span.TraceID == signal.TraceID:
        return Decision(True, 1.0, "trace_id_exact")
    if (span.Pod != "" and span.Pod == signal.Pod and span.PID > 0 and span.PID == signal.PID
            and within_window(span.Timestamp, signal.Timestamp, 100 * MS)):
        return Decision(True, 0.9, "pod_pid_100ms")
    if (span.Pod != "" and span.Pod == signal.Pod and span.ConnTuple != "" and span.ConnTuple == signal.ConnTuple
            and within_window(span.Timestamp, signal.Timestamp, 250 * MS)):
        return Decision(True, 0.8, "pod_conn_250ms")
    if (span.Service != "" and span.Service == signal.Service and span.Node != "" and span.Node == signal.Node
            and within_window(span.Timestamp, signal.Timestamp, 500 * MS)):
        return Decision(True, 0.65, "service_node_500ms")
    return Decision()


def EnrichDNS(base, span, signal, window, threshold):
    """dns.go:79-105"""
    if base is None:
        base = {}
    if threshold <= 0:
        threshold = DEFAULT_ENRICHMENT_THRESHOLD
    decision = Match(span, signal, window)
    if not decision.Matched or decision.Confidence < threshold:
        return base, decision
    if signal.Signal != "dns_latency_ms":
        return base, Decision()
    out = dict(base)
    out[ATTR["dns_latency_ms"]] = signal.Value
    out[ATTR_CORRELATION_CONF] = decision.Confidence
    return out, decision


# ---------------------------------------------------------------------------------------------
# pkg/correlation/evaluator.go
# ---------------------------------------------------------------------------------------------
@dataclass
class LabeledPair:  # evaluator.go:13-19
    CaseID: str
    Span: SpanRef
    Signal: SignalRef
    ExpectedMatch: bool
    ExpectedTier: str = ""


@dataclass
class Prediction:  # evaluator.go:22-31
    CaseID: str
    Expected: bool
    Predicted: bool
    Confidence: float
    Tier: str
    Correct: bool
    Signal: str
    ExpectedTier: str


@dataclass
class EvalReport:  # evaluator.go:34-51
    SampleSize: int = 0
    TruePositive: int = 0
    FalsePositive: int = 0
    FalseNegative: int = 0
    TrueNegative: int = 0
    Precision: float = 0.0
    Recall: float = 0.0
    F1: float = 0.0
    TierAccuracy: float = 0.0
    MeanConfidence: float = 0.0
    WindowMS: int = 0
    Threshold: float = 0.0


def safe_div(num: int, den: int) -> float:
    """evaluator.go:189-194"""
    if den == 0:
        return 0.0
    return float(num) / float(den)


def EvaluateLabeledPairs(pairs: List[LabeledPair], window: int, threshold: float) -> Tuple[EvalReport, List[Prediction]]:
    """evaluator.go:90-167"""
    if window <= 0:
        window = DEFAULT_WINDOW
    if threshold <= 0:
        threshold = DEFAULT_ENRICHMENT_THRESHOLD
    report = EvalReport(SampleSize=len(pairs), WindowMS=window // MS, Threshold=threshold)
    predictions: List[Prediction] = []
    tier_correct = tier_comparable = 0
    conf_sum = 0.0
    conf_count = 0
    for pair in pairs:
        decision = Match(pair.Span, pair.Signal, window)
        predicted = decision.Matched and decision.Confidence >= threshold
        correct = predicted == pair.ExpectedMatch
        predictions.append(Prediction(pair.CaseID, pair.ExpectedMatch, predicted, decision.Confidence, decision.Tier,
                                      correct, pair.Signal.Signal, pair.ExpectedTier))
        if predicted:
            conf_sum += decision.Confidence
            conf_count += 1
        if pair.ExpectedMatch and predicted:
            report.TruePositive += 1
        elif (not pair.ExpectedMatch) and predicted:
            report.FalsePositive += 1
        elif pair.ExpectedMatch and not predicted:
            report.FalseNegative += 1
        else:
            report.TrueNegative += 1
        if pair.ExpectedMatch and pair.ExpectedTier != "" and predicted:
            tier_comparable += 1
            if pair.ExpectedTier == decision.Tier:
                tier_correct += 1
    report.Precision = safe_div(report.TruePositive, report.TruePositive + report.FalsePositive)
    report.Recall = safe_div(report.TruePositive, report.TruePositive + report.FalseNegative)
    if report.Precision + report.Recall > 0:
        report.F1 = 2 * ((report.Precision * report.Recall) / (report.Precision + report.Recall))
    if tier_comparable > 0:
        report.TierAccuracy = float(tier_correct) / float(tier_comparable)
    if conf_count > 0:
        report.MeanConfidence = conf_sum / float(conf_count)
    report._tier_correct = tier_correct          # exposed for integer parity with lsg_eval
    report._tier_comparable = tier_comparable
    return report, predictions


def EvaluateGate(report: EvalReport, min_precision: float, min_recall: float) -> Tuple[bool, str]:
    """evaluator.go:170-187"""
    if report.Precision < min_precision:
        return False, "precision gate failed: got %.4f required %.4f" % (report.Precision, min_precision)
    if report.Recall < min_recall:
        return False, "recall gate failed: got %.4f required %.4f" % (report.Recall, min_recall)
    return True, "correlation gate passed"


# ---------------------------------------------------------------------------------------------
# pkg/correlation/retry_storm.go
# ---------------------------------------------------------------------------------------------
DEFAULT_STORM_WINDOW = 10 * SECOND
DEFAULT_STORM_THRESHOLD = 5


def _prune_old(events: List[int], now: int, window: int) -> List[int]:
    """retry_storm.go:98-110 (prefix-only pruning; assumes in-order arrival)"""
    cutoff = now - window
    start = 0
    while start < len(events) and events[start] < cutoff:
        start += 1
    return events[start:]


class RetryStormDetector:
    """retry_storm.go:21-96"""

    def __init__(self, window: int, threshold: int):
        self.window = window
        self.threshold = threshold
        self.buckets: Dict[str, List[int]] = {}

    def Record(self, pod: str, ts: int) -> bool:
        ev = self.buckets.setdefault(pod, [])
        ev.append(ts)
        self.buckets[pod] = _prune_old(ev, ts, self.window)
        return len(self.buckets[pod]) >= self.threshold

    def IsStorm(self, pod: str, now: int) -> bool:
        if pod not in self.buckets:
            return False
        self.buckets[pod] = _prune_old(self.buckets[pod], now, self.window)
        return len(self.buckets[pod]) >= self.threshold

    def Count(self, pod: str, now: int) -> int:
        if pod not in self.buckets:
            return 0
        self.buckets[pod] = _prune_old(self.buckets[pod], now, self.window)
        return len(self.buckets[pod])

    def Reset(self) -> None:
        self.buckets = {}


# ---------------------------------------------------------------------------------------------
# pkg/otel/processor/ebpfcorrelator
# ---------------------------------------------------------------------------------------------
@dataclass
class DebugStats:  # correlator.go:20-25
    Unmatched: int = 0
    LowConfidence: int = 0
    FanoutDropped: int = 0
    UnsupportedType: int = 0


@dataclass
class Candidate:  # correlator.go:28-31
    Signal: SignalRef
    Decision: Decision
    Index: int = -1  # position in the input signal slice (not in the reference struct; for parity checks)


@dataclass
class EnrichmentResult:  # correlator.go:34-38
    Attributes: Dict[str, float]
    Candidates: List[Candidate]
    Debug: DebugStats


def signal_attr_key(signal: str) -> Tuple[str, bool]:
    """correlator.go:143-172"""
    if signal in ATTR:
        return ATTR[signal], True
    return "", False


@dataclass
class Correlator:  # correlator.go:13-17, New() correlator.go:41-47
    WindowMS: int = 2000
    EnrichmentThreshold: float = 0.7
    MaxJoinFanout: int = 3

    def EnrichAttributes(self, base: Optional[Dict[str, float]], span: SpanRef, signals: List[SignalRef]) -> EnrichmentResult:
        """correlator.go:50-125"""
        if base is None:
            base = {}
        window = self.WindowMS * MS
        threshold = self.EnrichmentThreshold
        if threshold <= 0:
            threshold = DEFAULT_ENRICHMENT_THRESHOLD
        fanout = self.MaxJoinFanout
        if fanout <= 0:
            fanout = 3
        candidates: List[Candidate] = []
        out = dict(base)
        debug = DebugStats()
        for idx, signal in enumerate(signals):
            _, supported = signal_attr_key(signal.Signal)
            if not supported:
                debug.UnsupportedType += 1
                continue
            decision = Match(span, signal, window)
            if not decision.Matched:
                debug.Unmatched += 1
                continue
            if decision.Confidence < threshold:
                debug.LowConfidence += 1
                continue
            candidates.append(Candidate(signal, decision, idx))

        # sort.SliceStable, correlator.go:92-99: confidence desc, then |span.ts - sig.ts| asc; Python's sort is stable
        def key(c: Candidate):
            return (-c.Decision.Confidence, abs(float(_sub(span.Timestamp, c.Signal.Timestamp))))

        candidates.sort(key=key)
        if len(candidates) > fanout:
            debug.FanoutDropped = len(candidates) - fanout
            candidates = candidates[:fanout]
        max_conf = 0.0
        for c in candidates:
            attr, _ = signal_attr_key(c.Signal.Signal)
            if attr not in out or c.Signal.Value > out[attr]:
                out[attr] = c.Signal.Value
            if c.Decision.Confidence > max_conf:
                max_conf = c.Decision.Confidence
        if max_conf > 0:
            out[ATTR_CORRELATION_CONF] = max_conf
        return EnrichmentResult(out, candidates, debug)

    def EnrichDNSAttributes(self, base, span: SpanRef, signal: SignalRef):
        """correlator.go:128-141"""
        result = self.EnrichAttributes(base, span, [signal])
        if len(result.Candidates) == 0:
            return result.Attributes, Decision()
        return result.Attributes, result.Candidates[0].Decision

    def ProcessBatch(self, spans: List["SpanRecord"], signals: List[SignalRef]) -> "ProcessedBatch":
        """processor.go:29-49"""
        result = ProcessedBatch([], DebugStats())
        for item in spans:
            span_ref = SpanRef(item.TraceID, item.Service, item.Node, item.Pod, item.PID, item.ConnTuple, item.Timestamp)
            enriched = self.EnrichAttributes(item.Attributes, span_ref, signals)
            DecomposeRetrieval(enriched.Attributes)
            new_item = SpanRecord(item.TraceID, item.SpanID, item.Service, item.Node, item.Pod, item.PID,
                                  item.ConnTuple, item.Timestamp, enriched.Attributes)
            new_item._candidates = enriched.Candidates  # not in the reference struct; for parity checks
            result.Spans.append(new_item)
            result.Debug = merge_debug(result.Debug, enriched.Debug)
        return result


def New() -> Correlator:
    return Correlator()


def DecomposeRetrieval(attrs: Dict[str, float]) -> float:
    """correlator.go:179-194"""
    total = 0.0
    for key in (ATTR["dns_latency_ms"], ATTR["connect_latency_ms"], ATTR["tls_handshake_ms"]):
        if key in attrs:
            total += attrs[key]
    if total > 0:
        attrs[ATTR_RETRIEVAL_KERNEL_MS] = total
    return total


@dataclass
class SpanRecord:  # processor.go:10-20
    TraceID: str = ""
    SpanID: str = ""
    Service: str = ""
    Node: str = ""
    Pod: str = ""
    PID: int = 0
    ConnTuple: str = ""
    Timestamp: Optional[int] = None
    Attributes: Optional[Dict[str, float]] = None


@dataclass
class ProcessedBatch:  # processor.go:23-26
    Spans: List[SpanRecord]
    Debug: DebugStats


def merge_debug(left: DebugStats, right: DebugStats) -> DebugStats:
    """processor.go:51-58"""
    return DebugStats(left.Unmatched + right.Unmatched, left.LowConfidence + right.LowConfidence,
                      left.FanoutDropped + right.FanoutDropped, left.UnsupportedType + right.UnsupportedType)


# ---------------------------------------------------------------------------------------------
# pkg/attribution
# ---------------------------------------------------------------------------------------------
# bayesian.go:11-34
ALL_DOMAINS = ["network_dns", "network_egress", "cpu_throttle", "memory_pressure", "provider_throttle",
               "provider_error", "retrieval_backend", "unknown"]

# bayesian.go:69-189, rows in literal order, columns in AllDomains order
_LIK_ROWS = [
    ("dns_latency_ms", [0.95, 0.70, 0.10, 0.10, 0.10, 0.10, 0.15, 0.10]),
    ("tcp_retransmits_total", [0.15, 0.90, 0.10, 0.10, 0.10, 0.15, 0.10, 0.10]),
    ("runqueue_delay_ms", [0.10, 0.10, 0.90, 0.60, 0.10, 0.10, 0.10, 0.10]),
    ("connect_latency_ms", [0.50, 0.85, 0.10, 0.10, 0.75, 0.40, 0.30, 0.10]),
    ("tls_handshake_ms", [0.10, 0.30, 0.10, 0.10, 0.80, 0.50, 0.20, 0.10]),
    ("cpu_steal_pct", [0.10, 0.10, 0.90, 0.20, 0.10, 0.10, 0.10, 0.10]),
    ("cfs_throttled_ms", [0.10, 0.10, 0.85, 0.75, 0.10, 0.10, 0.10, 0.10]),
    ("mem_reclaim_latency_ms", [0.05, 0.05, 0.15, 0.95, 0.05, 0.05, 0.05, 0.05]),
    ("disk_io_latency_ms", [0.05, 0.05, 0.10, 0.85, 0.05, 0.05, 0.30, 0.05]),
    ("syscall_latency_ms", [0.10, 0.20, 0.15, 0.10, 0.90, 0.60, 0.40, 0.10]),
    ("connect_errors_total", [0.10, 0.80, 0.05, 0.05, 0.60, 0.85, 0.15, 0.10]),
    ("tls_handshake_fail_total", [0.05, 0.70, 0.05, 0.05, 0.30, 0.60, 0.10, 0.05]),
]
SIGNAL_ORDER = [name for name, _ in _LIK_ROWS]

# bayesian.go:193-206
SIGNAL_THRESHOLDS = {
    "dns_latency_ms": 40, "tcp_retransmits_total": 2, "runqueue_delay_ms": 10, "connect_latency_ms": 80,
    "tls_handshake_ms": 60, "cpu_steal_pct": 2, "cfs_throttled_ms": 40, "mem_reclaim_latency_ms": 5,
    "disk_io_latency_ms": 10, "syscall_latency_ms": 50, "connect_errors_total": 1, "tls_handshake_fail_total": 1,
}


def DefaultPriors() -> Dict[str, float]:
    """bayesian.go:54-62"""
    p = 1.0 / float(len(ALL_DOMAINS))
    return {d: p for d in ALL_DOMAINS}


def DefaultLikelihoods() -> Dict[str, Dict[str, float]]:
    """bayesian.go:67-190"""
    return {name: dict(zip(ALL_DOMAINS, row)) for name, row in _LIK_ROWS}


def clamp_likelihood(p: float) -> float:
    """bayesian.go:303-311"""
    if p < 0.01:
        return 0.01
    if p > 0.99:
        return 0.99
    return p


@dataclass
class Posterior:  # bayesian.go:210-214
    Domain: str
    Posterior: float
    Evidence: List[str]


class BayesianAttributor:
    """bayesian.go:39-52"""

    def __init__(self, priors=None, likelihoods=None):
        self.Priors = DefaultPriors() if priors is None else priors
        self.Likelihoods = DefaultLikelihoods() if likelihoods is None else likelihoods

    def likelihood_for(self, signal: str, domain: str, is_elevated: bool) -> float:
        """bayesian.go:288-301"""
        ll = self.Likelihoods.get(signal)
        if ll is None:
            return 0.5
        if domain not in ll:
            return 0.5
        p = ll[domain]
        if is_elevated:
            return clamp_likelihood(p)
        return clamp_likelihood(1 - p)

    def Attribute(self, signals: Dict[str, float]) -> List[Posterior]:
        """bayesian.go:218-284.  The reference iterates Go maps (random order); the order is fixed here:
        likelihood rows in dict insertion order (= literal order for the defaults), LSE in AllDomains order."""
        elevated = {}
        for signal, value in signals.items():
            if signal in SIGNAL_THRESHOLDS and value >= SIGNAL_THRESHOLDS[signal]:
                elevated[signal] = True
        log_post = {}
        for domain in ALL_DOMAINS:
            prior = self.Priors.get(domain, 0.0)
            if prior <= 0:
                prior = 1e-10
            log_p = math.log(prior)
            for signal in self.Likelihoods:
                log_p += math.log(self.likelihood_for(signal, domain, elevated.get(signal, False)))
            log_post[domain] = log_p
        max_log = -math.inf
        for d in ALL_DOMAINS:
            if log_post[d] > max_log:
                max_log = log_post[d]
        sum_exp = 0.0
        for d in ALL_DOMAINS:
            sum_exp += math.exp(log_post[d] - max_log)
        log_z = max_log + math.log(sum_exp)
        result = []
        for domain in ALL_DOMAINS:
            posterior = math.exp(log_post[domain] - log_z)
            evidence = sorted(s for s in elevated if s in self.Likelihoods and self.Likelihoods[s].get(domain, 0.0) >= 0.5)
            result.append(Posterior(domain, posterior, evidence))
        result.sort(key=lambda p: -p.Posterior)  # stable, bayesian.go:279-281
        return result

    def AttributeSample(self, sample: "FaultSample") -> "IncidentAttribution":
        """bayesian.go:315-343"""
        base = BuildAttribution(sample)
        if len(sample.Signals) == 0:
            return base
        posteriors = self.Attribute(sample.Signals)
        base.FaultHypotheses = [(p.Domain, p.Posterior, p.Evidence) for p in posteriors if not p.Posterior < 0.01]
        if posteriors:
            base.PredictedFaultDomain = posteriors[0].Domain
            base.Confidence = posteriors[0].Posterior
        return base


def NewBayesianAttributor() -> BayesianAttributor:
    return BayesianAttributor()


@dataclass
class FaultSample:  # mapper.go:11-26
    IncidentID: str = ""
    FaultLabel: str = ""
    ExpectedDomain: str = ""
    ExpectedDomains: List[str] = field(default_factory=list)
    Signals: Dict[str, float] = field(default_factory=dict)
    Confidence: float = 0.0


@dataclass
class IncidentAttribution:  # the fields of schema.IncidentAttribution this path writes
    IncidentID: str = ""
    PredictedFaultDomain: str = ""
    Confidence: float = 0.0
    FaultHypotheses: List[Tuple[str, float, List[str]]] = field(default_factory=list)


def MapFaultLabel(label: str) -> str:
    """mapper.go:29-50"""
    return {
        "dns_latency": "network_dns", "egress_drop": "network_egress", "cpu_throttle": "cpu_throttle",
        "memory_pressure": "memory_pressure", "network_partition": "network_egress",
        "provider_throttle": "provider_throttle", "provider_error": "provider_error",
        "retrieval_slowdown": "retrieval_backend",
    }.get(label, "unknown")


def BuildAttribution(sample: FaultSample) -> IncidentAttribution:
    """mapper.go:53-98 (numeric/label fields only; the evidence envelope is host-side string building)"""
    return IncidentAttribution(sample.IncidentID, MapFaultLabel(sample.FaultLabel), sample.Confidence, [])


def BuildAttributionsRule(samples):
    """pipeline.go:20-26"""
    return [BuildAttribution(s) for s in samples]


def BuildAttributionsBayes(samples, attributor=None):
    """pipeline.go:31-40"""
    if attributor is None:
        attributor = NewBayesianAttributor()
    return [attributor.AttributeSample(s) for s in samples]


def BuildAttributions(samples, mode: str):
    """pipeline.go:44-62"""
    if mode.strip().lower() == "rule":
        return BuildAttributionsRule(samples)
    return BuildAttributionsBayes(samples, None)


def _actual(sample: FaultSample) -> str:
    actual = sample.ExpectedDomain
    if actual == "":
        actual = MapFaultLabel(sample.FaultLabel)
    return actual


def BuildConfusionMatrix(samples, predictions) -> Dict[Tuple[str, str], int]:
    """pipeline.go:65-82"""
    matrix: Dict[Tuple[str, str], int] = {}
    for idx, prediction in enumerate(predictions):
        if idx >= len(samples):
            break
        key = (_actual(samples[idx]), prediction.PredictedFaultDomain)
        matrix[key] = matrix.get(key, 0) + 1
    return matrix


def Accuracy(samples, predictions) -> float:
    """pipeline.go:85-104"""
    if len(predictions) == 0:
        return 0.0
    correct = 0
    for idx, prediction in enumerate(predictions):
        if idx >= len(samples):
            break
        if _actual(samples[idx]) == prediction.PredictedFaultDomain:
            correct += 1
    return float(correct) / float(len(predictions))


def _expected_list(sample: FaultSample) -> List[str]:
    expected = sample.ExpectedDomains
    if len(expected) == 0:
        expected = [_actual(sample)]
    return expected


def PartialAccuracy(samples, predictions) -> float:
    """pipeline.go:110-136"""
    if len(predictions) == 0:
        return 0.0
    correct = 0
    for idx, prediction in enumerate(predictions):
        if idx >= len(samples):
            break
        for domain in _expected_list(samples[idx]):
            if domain == prediction.PredictedFaultDomain:
                correct += 1
                break
    return float(correct) / float(len(predictions))


def CoverageAccuracy(samples, predictions, threshold: float) -> float:
    """pipeline.go:140-185"""
    if len(predictions) == 0:
        return 0.0
    total = 0.0
    count = 0
    for idx, prediction in enumerate(predictions):
        if idx >= len(samples):
            break
        expected = _expected_list(samples[idx])
        hyp = {}
        for (domain, posterior, _e) in prediction.FaultHypotheses:
            if posterior >= threshold:
                hyp[domain] = True
        hyp[prediction.PredictedFaultDomain] = True
        found = 0
        for domain in expected:
            if hyp.get(domain, False):
                found += 1
        total += float(found) / float(len(expected))
        count += 1
    if count == 0:
        return 0.0
    return total / float(count)


# --------------------------------------------------------------------------------------------------
# pkg/slo/calculator.go (quantile; pkg/releasegate/gate.go:737-760 has the same body)
def quantile(values, q: float) -> float:
    """calculator.go:121-149"""
    if len(values) == 0:
        return 0.0
    if q <= 0:
        q = 0.0
    if q >= 1:
        q = 1.0
    # sort.Float64s: ascending, NaN first
    sorted_ = sorted(values, key=lambda v: (0, 0.0) if v != v else (1, v))
    if len(sorted_) == 1:
        return sorted_[0]
    pos = q * float(len(sorted_) - 1)
    lower = int(math.floor(pos))
    upper = int(math.ceil(pos))
    if lower == upper:
        return sorted_[lower]
    frac = pos - float(lower)
    return sorted_[lower] * (1 - frac) + sorted_[upper] * frac
