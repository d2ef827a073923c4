"""Mirror of pkg/correlation (dns.go, evaluator.go) over the GPU path.

Same names, argument meaning and defaults as the Go package; the arithmetic (Match tiers, window
tests, TP/FP/FN/TN counting) runs in libllmslo_gpu.so's K-pairs kernel.
"""
import json
from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from . import _abi as A
from ._lib import default_context
from .interner import Interner

MS = 1_000_000
SECOND = 1_000_000_000
DefaultWindow = 2 * SECOND            # dns.go:12
DefaultEnrichmentThreshold = 0.7      # dns.go:15


@dataclass
class SpanRef:  # dns.go:19-27
    TraceID: str = ""
    Service: str = ""
    Node: str = ""
    Pod: str = ""
    PID: int = 0
    ConnTuple: str = ""
    Timestamp: Optional[int] = None   # Unix ns, datetime, or None for the zero Time


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


def _decision(tier):
    if tier == 0:
        return Decision()
    return Decision(True, A.TIER_CONF[tier], A.TIER_NAMES[tier])


def MatchBatch(spans: List[SpanRef], signals: List[SignalRef], window: int, ctx=None) -> List[Decision]:
    """Match() for n independent (span, signal) pairs in one kernel launch."""
    if len(spans) == 0:
        return []
    ctx = ctx or default_context()
    it = Interner()
    sp, ev = it.span_records(spans), it.event_records(signals)
    tier, _pred, _e = ctx.match_pairs(sp, ev, it.ident_table(), window, 0.0)
    return [_decision(int(t)) for t in tier]


def Match(span: SpanRef, signal: SignalRef, window: int, ctx=None) -> Decision:
    """dns.go:50-76"""
    return MatchBatch([span], [signal], window, ctx)[0]


def EnrichDNS(base, span, signal, window, threshold, ctx=None):
    """dns.go:79-105"""
    if base is None:
        base = {}
    if threshold <= 0:
        threshold = DefaultEnrichmentThreshold
    decision = Match(span, signal, window, ctx)
    if not decision.Matched or decision.Confidence < threshold:
        return base, decision
    if signal.Signal != "dns_latency_ms":
        return base, Decision()
    out = dict(base)
    out["llm.ebpf.dns.latency_ms"] = signal.Value
    out["llm.ebpf.correlation_confidence"] = decision.Confidence
    return out, decision


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
    MinPrecisionReq: float = 0.0
    MinRecallReq: float = 0.0
    PassedGate: bool = False


@dataclass
class GateResult:  # evaluator.go:54-57
    Pass: bool
    Message: str


def _parse_time(s):
    from datetime import datetime, timezone
    import re
    if not s or s.startswith("0001-01-01T00:00:00"):
        return None
    m = re.match(r"^(\d{4})-(\d\d)-(\d\d)T(\d\d):(\d\d):(\d\d)(?:\.(\d{1,9}))?Z$", s)
    if not m:
        raise ValueError("parse labeled pair: unsupported timestamp %r" % s)
    y, mo, d, h, mi, sec, frac = m.groups()
    base = int(datetime(int(y), int(mo), int(d), int(h), int(mi), int(sec), tzinfo=timezone.utc).timestamp())
    return base * SECOND + int((frac or "0").ljust(9, "0"))


def LoadLabeledPairsFromJSONL(path: str) -> List[LabeledPair]:
    """evaluator.go:60-87; errors are raised as OSError / ValueError with the reference's messages."""
    try:
        f = open(path)
    except OSError as e:
        raise OSError("open labeled pairs file: %s" % e) from e
    out = []
    with f:
        for line in f:
            line = line.strip()
            if not line:
                continue
            try:
                r = json.loads(line)
            except ValueError as e:
                raise ValueError("parse labeled pair: %s" % e) from e
            sp, sg = r.get("span", {}), r.get("signal", {})
            out.append(LabeledPair(
                r.get("case_id", ""),
                SpanRef(sp.get("trace_id", ""), sp.get("service", ""), sp.get("node", ""), sp.get("pod", ""),
                        int(sp.get("pid", 0)), sp.get("conn_tuple", ""), _parse_time(sp.get("timestamp"))),
                SignalRef(sg.get("signal", ""), sg.get("trace_id", ""), sg.get("service", ""), sg.get("node", ""),
                          sg.get("pod", ""), int(sg.get("pid", 0)), sg.get("conn_tuple", ""),
                          _parse_time(sg.get("timestamp")), float(sg.get("value", 0.0))),
                bool(r.get("expected_match", False)), r.get("expected_tier", "")))
    if not out:
        raise ValueError("no labeled pairs loaded from %s" % path)
    return out


_TIER_CODE = {name: i for i, name in enumerate(A.TIER_NAMES) if name}


def _safe_div(num, den):  # evaluator.go:189-194
    return 0.0 if den == 0 else float(num) / float(den)


def EvaluateLabeledPairs(pairs: List[LabeledPair], window: int, threshold: float, ctx=None):
    """evaluator.go:90-167.  Per-pair decisions and the integer counters come from the GPU; the five
    ratios are formed here from those integers."""
    if window <= 0:
        window = DefaultWindow
    if threshold <= 0:
        threshold = DefaultEnrichmentThreshold
    report = EvalReport(SampleSize=len(pairs), WindowMS=window // MS, Threshold=threshold)
    if len(pairs) == 0:
        return report, []
    ctx = ctx or default_context()
    it = Interner()
    sp = it.span_records([p.Span for p in pairs])
    ev = it.event_records([p.Signal for p in pairs])
    em = np.array([p.ExpectedMatch for p in pairs], np.uint8)
    # an expected tier string outside the four known names can never equal decision.Tier: code 7
    et = np.array([0 if p.ExpectedTier == "" else _TIER_CODE.get(p.ExpectedTier, 7) for p in pairs], np.uint8)
    tier, pred, e = ctx.match_pairs(sp, ev, it.ident_table(), window, threshold, em, et)
    predictions = []
    for p, t, pr in zip(pairs, tier, pred):
        d = _decision(int(t))
        predictions.append(Prediction(p.CaseID, p.ExpectedMatch, bool(pr), d.Confidence, d.Tier,
                                      bool(pr) == p.ExpectedMatch, p.Signal.Signal, p.ExpectedTier))
    report.TruePositive, report.FalsePositive = int(e.tp), int(e.fp)
    report.FalseNegative, report.TrueNegative = int(e.fn), int(e.tn)
    report.Precision = _safe_div(report.TruePositive, report.TruePositive + report.FalsePositive)
    report.Recall = _safe_div(report.TruePositive, report.TruePositive + report.FalseNegative)
    if report.Precision + report.Recall > 0:
        report.F1 = 2 * ((report.Precision * report.Recall) / (report.Precision + report.Recall))
    if e.tier_comparable > 0:
        report.TierAccuracy = float(e.tier_correct) / float(e.tier_comparable)
    n_pred = sum(int(x) for x in e.predicted_by_tier)
    if n_pred > 0:
        report.MeanConfidence = sum(int(e.predicted_by_tier[t]) * A.TIER_CONF[t] for t in range(5)) / float(n_pred)
    return report, predictions


def EvaluateGate(report: EvalReport, minPrecision: float, minRecall: float) -> GateResult:
    """evaluator.go:170-187"""
    if report.Precision < minPrecision:
        return GateResult(False, "precision gate failed: got %.4f required %.4f" % (report.Precision, minPrecision))
    if report.Recall < minRecall:
        return GateResult(False, "recall gate failed: got %.4f required %.4f" % (report.Recall, minRecall))
    return GateResult(True, "correlation gate passed")


# ---- pkg/correlation/retry_storm.go -------------------------------------------------------------------------
DefaultStormWindow = 10 * SECOND      # retry_storm.go:9-10
DefaultStormThreshold = 5             # retry_storm.go:12-14


class RetryStormDetector:
    """retry_storm.go:21-96.  RecordBatch is the GPU path: one sort by (pod, time) and one window count for the
    whole batch (lsg_retry_storm).  Record / IsStorm / Count keep the reference's streaming interface on top of
    it: the detector retains, per pod, the timestamps still inside the window (what the reference's bucket holds
    after pruneOld) and prepends them to the next batch."""

    def __init__(self, window: int, threshold: int, ctx=None):
        self.window, self.threshold, self.ctx = window, threshold, ctx
        self.buckets = {}

    def RecordBatch(self, pods, timestamps):
        """Record() for every (pod, ts) in arrival order; returns the list of Record's return values."""
        if len(pods) == 0:
            return []
        ctx = self.ctx or default_context()
        from .interner import to_ns
        pod_ids = {}
        hist_p, hist_t = [], []
        for pod, ts_list in self.buckets.items():
            for t in ts_list:
                hist_p.append(pod)
                hist_t.append(t)
        all_p = hist_p + list(pods)
        all_t = hist_t + [to_ns(t) for t in timestamps]
        ev = np.zeros(len(all_p), A.EVENT_DTYPE)
        for i, (pod, t) in enumerate(zip(all_p, all_t)):
            ev[i] = (t, 1.0, pod_ids.setdefault(pod, len(pod_ids)), 0, 0, A.SIGNAL_INDEX["tcp_retransmits_total"])
        idents = np.zeros(max(len(pod_ids), 1), A.IDENT_DTYPE)
        for pod, row in pod_ids.items():
            idents[row] = (row + 1, 0, 0, 1)
        cnt, _storm = ctx.retry_storm(ev, idents, self.window, self.threshold)
        out = [int(c) >= self.threshold for c in cnt[len(hist_p):]]
        # what stays in each bucket: the events of the pod not older than its newest event minus the window
        newest = {}
        for pod, t in zip(all_p, all_t):
            newest[pod] = t if pod not in newest else max(newest[pod], t)
        self.buckets = {}
        for pod, t in zip(all_p, all_t):
            if t >= newest[pod] - self.window:
                self.buckets.setdefault(pod, []).append(t)
        return out

    def Record(self, pod: str, ts) -> bool:
        """retry_storm.go:46-62"""
        return self.RecordBatch([pod], [ts])[0]

    def Count(self, pod: str, now) -> int:
        """retry_storm.go:80-90"""
        from .interner import to_ns
        if pod not in self.buckets:
            return 0
        now = to_ns(now)
        self.buckets[pod] = [t for t in self.buckets[pod] if t >= now - self.window]
        return len(self.buckets[pod])

    def IsStorm(self, pod: str, now) -> bool:
        """retry_storm.go:66-77"""
        return pod in self.buckets and self.Count(pod, now) >= self.threshold

    def Reset(self) -> None:
        """retry_storm.go:93-96"""
        self.buckets = {}


def NewRetryStormDetector(window: int, threshold: int, ctx=None) -> RetryStormDetector:
    return RetryStormDetector(window, threshold, ctx)
