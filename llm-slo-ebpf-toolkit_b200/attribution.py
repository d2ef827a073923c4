"""Mirror of pkg/attribution (bayesian.go, pipeline.go, mapper.go, io.go) over the GPU path.

Posteriors, elevated masks, argmax labels, the confusion matrix and the accuracy numerators come
from libllmslo_gpu.so (K-bayes, K-confusion).  Evidence lists, the >= 0.01 hypothesis filter, the
stable ordering of eight posteriors and envelope building are host-side string work, as the C ABI
assigns them.
"""
import ctypes as C
import json
import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

from . import _abi as A
from ._lib import default_context, load_library

AllDomainsList = list(A.DOMAINS)
AttributionModeBayes = "bayes"
AttributionModeRule = "rule"


def AllDomains() -> List[str]:
    return list(A.DOMAINS)


def DefaultPriors() -> Dict[str, float]:
    """bayesian.go:54-62"""
    m = A.Model()
    load_library().lsg_default_model(C.byref(m))
    return {d: m.prior[i] for i, d in enumerate(A.DOMAINS)}


def DefaultLikelihoods() -> Dict[str, Dict[str, float]]:
    """bayesian.go:67-190"""
    m = A.Model()
    load_library().lsg_default_model(C.byref(m))
    return {s: {d: m.lik[i][j] for j, d in enumerate(A.DOMAINS)} for i, s in enumerate(A.SIGNALS)}


def _thresholds():
    m = A.Model()
    load_library().lsg_default_model(C.byref(m))
    return [m.thresh[i] for i in range(A.NSIG)]


@dataclass
class Posterior:  # bayesian.go:210-214
    Domain: str
    Posterior: float
    Evidence: List[str]


@dataclass
class FaultSample:  # mapper.go:11-26
    IncidentID: str = ""
    Timestamp: Optional[str] = None
    Cluster: str = ""
    Namespace: str = ""
    Service: str = ""
    FaultLabel: str = ""
    ExpectedDomain: str = ""
    ExpectedDomains: List[str] = field(default_factory=list)
    Signals: Dict[str, float] = field(default_factory=dict)
    Confidence: float = 0.0
    BurnRate: float = 0.0
    WindowMinutes: int = 0
    RequestID: str = ""
    TraceID: str = ""


@dataclass
class FaultHypothesis:  # schema.FaultHypothesis
    Domain: str
    Posterior: float
    Evidence: List[str]


@dataclass
class IncidentAttribution:  # schema.IncidentAttribution (pkg/schema/types.go)
    IncidentID: str = ""
    Timestamp: Optional[str] = None
    Cluster: str = ""
    Namespace: str = ""
    Service: str = ""
    PredictedFaultDomain: str = ""
    Confidence: float = 0.0
    Evidence: List[dict] = field(default_factory=list)
    SLOImpact: dict = field(default_factory=dict)
    TraceIDs: List[str] = field(default_factory=list)
    RequestIDs: List[str] = field(default_factory=list)
    FaultHypotheses: List[FaultHypothesis] = field(default_factory=list)


_LABEL_MAP = {"dns_latency": "network_dns", "egress_drop": "network_egress", "cpu_throttle": "cpu_throttle",
              "memory_pressure": "memory_pressure", "network_partition": "network_egress",
              "provider_throttle": "provider_throttle", "provider_error": "provider_error",
              "retrieval_slowdown": "retrieval_backend"}


def MapFaultLabel(label: str) -> str:
    """mapper.go:29-50 (a string table lookup; stays on the host like all string work)"""
    return _LABEL_MAP.get(label, "unknown")


def BuildAttribution(sample: FaultSample) -> IncidentAttribution:
    """mapper.go:53-98"""
    domain = MapFaultLabel(sample.FaultLabel)
    evidence = [{"signal": "fault_label", "value": sample.FaultLabel, "source": "application"},
                {"signal": "mapped_domain", "value": domain, "source": "ebpf"},
                {"signal": "llm.ebpf.correlation_confidence", "value": sample.Confidence, "source": "otel"}]
    if sample.FaultLabel == "dns_latency":
        evidence.append({"signal": "llm.ebpf.dns.latency_ms", "value": 180.0, "source": "ebpf"})
    return IncidentAttribution(sample.IncidentID, sample.Timestamp, sample.Cluster, sample.Namespace, sample.Service,
                               domain, sample.Confidence, evidence,
                               {"sli": "ttft_ms", "burn_rate": sample.BurnRate, "window_minutes": sample.WindowMinutes},
                               [sample.TraceID], [sample.RequestID])


class BayesianAttributor:
    """bayesian.go:39-52.  Priors and Likelihoods are public and mutable as in Go; they are flattened
    into an lsg_model on every call, so edits take effect."""

    def __init__(self, ctx=None):
        self.Priors = DefaultPriors()
        self.Likelihoods = DefaultLikelihoods()
        self.ctx = ctx

    def _model(self):
        m = A.Model()
        th = _thresholds()
        mask = 0
        for d, name in enumerate(A.DOMAINS):
            m.prior[d] = float(self.Priors.get(name, 0.0))
            m.log_extra[d] = 0.0
        for s, name in enumerate(A.SIGNALS):
            m.thresh[s] = th[s]
            ll = self.Likelihoods.get(name)
            if ll is not None:
                mask |= 1 << s
            for d, dn in enumerate(A.DOMAINS):
                m.lik[s][d] = float(ll[dn]) if (ll is not None and dn in ll) else math.nan
        # Likelihoods rows outside the 12 known signals can never be elevated (no threshold,
        # bayesian.go:222) and add a constant log clamp(1-p) per domain
        for name, ll in self.Likelihoods.items():
            if name in A.SIGNAL_INDEX:
                continue
            for d, dn in enumerate(A.DOMAINS):
                p = 1.0 - ll[dn] if dn in ll else 0.5
                if dn in ll:
                    p = min(max(p, 0.01), 0.99)
                m.log_extra[d] += math.log(p)
        m.sig_mask = mask
        return m

    def _run(self, sig_maps: List[Dict[str, float]], rule_labels=None):
        ctx = self.ctx or default_context()
        ctx.set_model(self._model())
        n = len(sig_maps)
        arr = np.full((n, A.NSIG), np.nan)
        has = np.zeros(n, np.uint8)
        for i, sm in enumerate(sig_maps):
            has[i] = 1 if len(sm) > 0 else 0
            for k, v in sm.items():
                j = A.SIGNAL_INDEX.get(k)
                if j is not None:
                    arr[i, j] = v
        return ctx.attribute(arr, has, rule_labels)

    def _posteriors(self, post_row, elev) -> List[Posterior]:
        result = []
        for d, dn in enumerate(A.DOMAINS):
            ev = sorted(A.SIGNALS[s] for s in range(A.NSIG)
                        if (elev >> s) & 1 and A.SIGNALS[s] in self.Likelihoods
                        and self.Likelihoods[A.SIGNALS[s]].get(dn, 0.0) >= 0.5)
            result.append(Posterior(dn, float(post_row[d]), ev))
        result.sort(key=lambda p: -p.Posterior)   # stable, bayesian.go:279-281
        return result

    def Attribute(self, signals: Dict[str, float]) -> List[Posterior]:
        """bayesian.go:218-284.  Unlike AttributeSample, Attribute itself has no empty-map shortcut:
        an empty map means no signal is elevated."""
        ctx = self.ctx or default_context()
        ctx.set_model(self._model())
        arr = np.full((1, A.NSIG), np.nan)
        for k, v in signals.items():
            j = A.SIGNAL_INDEX.get(k)
            if j is not None:
                arr[0, j] = v
        post, _top1, elev = ctx.attribute(arr, np.ones(1, np.uint8), None)
        return self._posteriors(post[0], int(elev[0]))

    def AttributeSamples(self, samples: List[FaultSample]) -> List[IncidentAttribution]:
        """AttributeSample (bayesian.go:315-343) for a batch: one kernel launch for all incidents."""
        if len(samples) == 0:
            return []
        rule = np.array([A.DOMAIN_INDEX[MapFaultLabel(s.FaultLabel)] for s in samples], np.uint8)
        post, top1, elev = self._run([s.Signals or {} for s in samples], rule)
        out = []
        for i, s in enumerate(samples):
            base = BuildAttribution(s)
            if len(s.Signals or {}) == 0:
                out.append(base)
                continue
            posteriors = self._posteriors(post[i], int(elev[i]))
            base.FaultHypotheses = [FaultHypothesis(p.Domain, p.Posterior, p.Evidence) for p in posteriors
                                    if not p.Posterior < 0.01]
            base.PredictedFaultDomain = A.DOMAINS[int(top1[i])]
            base.Confidence = float(post[i][int(top1[i])])
            out.append(base)
        return out

    def AttributeSample(self, sample: FaultSample) -> IncidentAttribution:
        """bayesian.go:315-343"""
        return self.AttributeSamples([sample])[0]


def NewBayesianAttributor(ctx=None) -> BayesianAttributor:
    return BayesianAttributor(ctx)


def BuildAttributionsRule(samples):
    """pipeline.go:20-26"""
    return [BuildAttribution(s) for s in samples]


def BuildAttributionsBayes(samples, attributor=None):
    """pipeline.go:31-40"""
    if attributor is None:
        attributor = NewBayesianAttributor()
    return attributor.AttributeSamples(list(samples))


def BuildAttributions(samples, mode: str):
    """pipeline.go:44-62"""
    if mode.strip().lower() == AttributionModeRule:
        return BuildAttributionsRule(samples)
    return BuildAttributionsBayes(samples, None)


@dataclass(frozen=True)
class MatrixKey:  # pipeline.go:8-11
    Actual: str
    Predicted: str


def _encode(samples, predictions, threshold):
    """FaultSample / IncidentAttribution lists -> the integer columns lsg_confusion takes."""
    n = min(len(samples), len(predictions))
    classes = {d: i for i, d in enumerate(A.DOMAINS)}
    actual = np.zeros(n, np.uint8)
    pred = np.zeros(n, np.uint8)
    exp = np.zeros(n, np.uint32)
    exp_n = np.zeros(n, np.uint8)
    post = np.full((n, A.NDOM), np.nan)
    other = {}
    for i in range(n):
        s, p = samples[i], predictions[i]
        a = s.ExpectedDomain or MapFaultLabel(s.FaultLabel)
        if a in classes:
            actual[i] = classes[a]
        else:
            if a not in other:
                if len(other) >= A.MAX_ACTUAL - A.NDOM:
                    raise ValueError("more than %d actual classes outside AllDomains" % (A.MAX_ACTUAL - A.NDOM))
                other[a] = A.NDOM + len(other)
            actual[i] = other[a]
        if p.PredictedFaultDomain not in classes:
            raise ValueError("predicted domain %r is not one of AllDomains" % p.PredictedFaultDomain)
        pred[i] = classes[p.PredictedFaultDomain]
        expected = list(s.ExpectedDomains) if s.ExpectedDomains else [a]
        if len(expected) > 8:
            raise ValueError("more than 8 expected domains")
        code = 0
        for k, dname in enumerate(expected):
            code |= (classes.get(dname, 0xF)) << (4 * k)
        exp[i], exp_n[i] = code, len(expected)
        for h in p.FaultHypotheses:
            post[i, classes[h.Domain]] = h.Posterior
    names = list(A.DOMAINS) + [None] * (A.MAX_ACTUAL - A.NDOM)
    for k, v in other.items():
        names[v] = k
    return n, actual, pred, exp, exp_n, post, names


def _confusion(samples, predictions, threshold=0.0, ctx=None):
    ctx = ctx or default_context()
    n, actual, pred, exp, exp_n, post, names = _encode(samples, predictions, threshold)
    return ctx.confusion(actual, pred, exp, exp_n, post, threshold), names, n


def BuildConfusionMatrix(samples, predictions, ctx=None) -> Dict[MatrixKey, int]:
    """pipeline.go:65-82"""
    if min(len(samples), len(predictions)) == 0:
        return {}
    out, names, _ = _confusion(samples, predictions, 0.0, ctx)
    m = {}
    for a in range(A.MAX_ACTUAL):
        for p in range(A.NDOM):
            c = int(out.matrix[a][p])
            if c:
                m[MatrixKey(names[a], A.DOMAINS[p])] = c
    return m


def Accuracy(samples, predictions, ctx=None) -> float:
    """pipeline.go:85-104 (the denominator is len(predictions), as in Go)"""
    if len(predictions) == 0:
        return 0.0
    if len(samples) == 0:
        return 0.0
    out, _, _ = _confusion(samples, predictions, 0.0, ctx)
    return float(out.correct) / float(len(predictions))


def PartialAccuracy(samples, predictions, ctx=None) -> float:
    """pipeline.go:110-136"""
    if len(predictions) == 0 or len(samples) == 0:
        return 0.0
    out, _, _ = _confusion(samples, predictions, 0.0, ctx)
    return float(out.partial) / float(len(predictions))


def CoverageAccuracy(samples, predictions, threshold: float, ctx=None) -> float:
    """pipeline.go:140-185: mean over incidents of found/len(expected), formed from the GPU's
    (found, len) histogram."""
    if len(predictions) == 0 or len(samples) == 0:
        return 0.0
    out, _, n = _confusion(samples, predictions, threshold, ctx)
    total = 0.0
    for f in range(A.NDOM + 1):
        for ln in range(1, A.NDOM + 1):
            c = int(out.coverage[f][ln])
            if c:
                total += c * (float(f) / float(ln))
    return total / float(n) if n else 0.0


def LoadSamplesFromJSONL(path: str) -> List[FaultSample]:
    """io.go:12-39"""
    try:
        f = open(path)
    except OSError as e:
        raise OSError("open samples file: %s" % e) from e
    out = []
    with f:
        for line in f:
            line = line.strip()
            if not line:
                continue
            try:
                r = json.loads(line)
            except ValueError as e:
                raise ValueError("parse sample: %s" % e) from e
            out.append(FaultSample(r.get("incident_id", ""), r.get("timestamp"), r.get("cluster", ""),
                                   r.get("namespace", ""), r.get("service", ""), r.get("fault_label", ""),
                                   r.get("expected_domain", ""), list(r.get("expected_domains") or []),
                                   dict(r.get("signals") or {}), float(r.get("confidence", 0.0)),
                                   float(r.get("burn_rate", 0.0)), int(r.get("window_minutes", 0)),
                                   r.get("request_id", ""), r.get("trace_id", "")))
    if not out:
        raise ValueError("no samples loaded from %s" % path)
    return out
