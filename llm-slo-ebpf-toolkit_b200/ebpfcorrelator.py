"""Mirror of pkg/otel/processor/ebpfcorrelator (correlator.go, processor.go) over the GPU path.

ProcessBatch is ONE C-ABI call: sort events by (join group, time), windowed join per span,
top-`fanout` selection, per-attribute max, confidence, DecomposeRetrieval and the four debug
counters all come back from libllmslo_gpu.so.  This module only interns strings on the way in and
turns attribute slots back into the semconv-keyed maps on the way out.
"""
import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

from . import _abi as A
from ._lib import default_context
from .correlation import MS, Decision, SignalRef, SpanRef, _decision
from .interner import Interner

# pkg/semconv/llm_ebpf.go, in signal-index order (signalAttrKey, correlator.go:143-172)
ATTR_KEYS = ["llm.ebpf.dns.latency_ms", "llm.ebpf.tcp.retransmits", "llm.ebpf.sched.runqueue_delay_ms",
             "llm.ebpf.net.connect_latency_ms", "llm.ebpf.tls.handshake_ms", "llm.ebpf.cpu.steal_pct",
             "llm.ebpf.cpu.cfs_throttled_ms", "llm.ebpf.mm.reclaim_latency_ms", "llm.ebpf.blk.io_latency_ms",
             "llm.ebpf.syscall.latency_ms", "llm.ebpf.net.connect_errors_total", "llm.ebpf.tls.handshake_fail_total"]
ATTR_SLOT = {k: i for i, k in enumerate(ATTR_KEYS)}
AttrCorrelationConf = "llm.ebpf.correlation_confidence"
AttrRetrievalKernelMS = "llm.ebpf.retrieval.kernel_attributed_ms"


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
    Index: int = -1


@dataclass
class EnrichmentResult:  # correlator.go:34-38
    Attributes: Dict[str, float]
    Candidates: List[Candidate]
    Debug: DebugStats


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
    Candidates: List[Candidate] = field(default_factory=list)   # not in the Go struct: kept for diagnostics


@dataclass
class ProcessedBatch:  # processor.go:23-26
    Spans: List[SpanRecord]
    Debug: DebugStats


def _base_slots(bases):
    """Attribute maps -> [n, 12] slots, NaN = absent.  Keys outside the 12 slots are carried on the host."""
    arr = np.full((len(bases), A.NSIG), np.nan)
    any_base = False
    for i, b in enumerate(bases):
        if b:
            for k, v in b.items():
                j = ATTR_SLOT.get(k)
                if j is not None:
                    arr[i, j] = v
                    any_base = True
    return arr if any_base else None


@dataclass
class Correlator:  # correlator.go:13-17
    WindowMS: int = 2000
    EnrichmentThreshold: float = 0.7
    MaxJoinFanout: int = 3
    ctx: object = None

    def _run(self, span_refs, bases, signals, decompose):
        ctx = self.ctx or default_context()
        it = Interner()
        sp = it.span_records(span_refs)
        ev = it.event_records(signals)
        fanout = self.MaxJoinFanout if self.MaxJoinFanout > 0 else 3
        # Go: time.Duration(c.WindowMS) * time.Millisecond, passed to Match where <= 0 means the default
        res = ctx.process_batch(sp, ev, it.ident_table(), self.WindowMS * MS, self.EnrichmentThreshold, fanout,
                                _base_slots(bases))
        outs = []
        for i in range(len(span_refs)):
            attrs = dict(bases[i]) if bases[i] else {}
            row = res["attr"][i]
            for j in range(A.NSIG):
                if not math.isnan(row[j]):
                    attrs[ATTR_KEYS[j]] = float(row[j])
            if res["conf"][i] > 0:
                attrs[AttrCorrelationConf] = float(res["conf"][i])
            if decompose and res["retrieval"][i] > 0:
                attrs[AttrRetrievalKernelMS] = float(res["retrieval"][i])
            cands = []
            for k in range(fanout):
                e = int(res["cand_event"][i, k])
                if e != A.NONE:
                    cands.append(Candidate(signals[e], _decision(int(res["cand_tier"][i, k])), e))
            outs.append((attrs, cands))
        d = res["debug"]
        return outs, DebugStats(int(d.unmatched), int(d.low_confidence), int(d.fanout_dropped),
                                int(d.unsupported_type)), res

    def EnrichAttributes(self, base, span: SpanRef, signals: List[SignalRef]) -> EnrichmentResult:
        """correlator.go:50-125"""
        outs, dbg, _ = self._run([span], [base], list(signals), decompose=False)
        return EnrichmentResult(outs[0][0], outs[0][1], dbg)

    def EnrichDNSAttributes(self, base, span: SpanRef, signal: SignalRef):
        """correlator.go:128-141"""
        result = self.EnrichAttributes(base, span, [signal])
        if len(result.Candidates) == 0:
            return result.Attributes, Decision()
        return result.Attributes, result.Candidates[0].Decision

    def ProcessBatch(self, spans: List[SpanRecord], signals: List[SignalRef]) -> ProcessedBatch:
        """processor.go:29-49"""
        if len(spans) == 0:
            return ProcessedBatch([], DebugStats())
        refs = [SpanRef(s.TraceID, s.Service, s.Node, s.Pod, s.PID, s.ConnTuple, s.Timestamp) for s in spans]
        outs, dbg, _ = self._run(refs, [s.Attributes for s in spans], list(signals), decompose=True)
        out_spans = [SpanRecord(s.TraceID, s.SpanID, s.Service, s.Node, s.Pod, s.PID, s.ConnTuple, s.Timestamp, a, c)
                     for s, (a, c) in zip(spans, outs)]
        return ProcessedBatch(out_spans, dbg)


def New(ctx=None) -> Correlator:
    """correlator.go:41-47"""
    return Correlator(2000, 0.7, 3, ctx)


def DecomposeRetrieval(attrs: Dict[str, float]) -> float:
    """correlator.go:179-194 for one attribute map handed in by the caller.  Inside ProcessBatch the
    sum is computed on the GPU; this standalone form is three additions on a caller-owned map and is
    kept for API completeness (the reference exports it)."""
    total = 0.0
    for key in (ATTR_KEYS[0], ATTR_KEYS[3], ATTR_KEYS[4]):
        if key in attrs:
            total += attrs[key]
    if total > 0:
        attrs[AttrRetrievalKernelMS] = total
    return total
