//go:build gpu

package ebpfcorrelator

import (
	"math"
	"time"

	"github.com/ogulcanaydogan/llm-slo-ebpf-toolkit/pkg/correlation"
	"github.com/ogulcanaydogan/llm-slo-ebpf-toolkit/pkg/gpu"
	"github.com/ogulcanaydogan/llm-slo-ebpf-toolkit/pkg/semconv"
)

// attribute key per signal slot (signalAttrKey, correlator.go:143-172), in gpu.Signals order
var attrKeys = [...]string{semconv.AttrDNSLatencyMS, semconv.AttrTCPRetransmits, semconv.AttrRunqueueDelayMS,
	semconv.AttrConnectLatencyMS, semconv.AttrTLSHandshakeMS, semconv.AttrCPUStealPct, semconv.AttrCFSThrottledMS,
	semconv.AttrMemReclaimLatencyMS, semconv.AttrDiskIOLatencyMS, semconv.AttrSyscallLatencyMS,
	semconv.AttrConnectErrors, semconv.AttrTLSHandshakeFails}

func (c Correlator) params() (int64, float64, int) {
	return int64(time.Duration(c.WindowMS) * time.Millisecond), c.EnrichmentThreshold, c.MaxJoinFanout // defaults applied by the library
}

// enrichBatch packs spans and signals, runs lsg_process_batch and rebuilds what EnrichAttributes returns per span.
func (c Correlator) enrichBatch(bases []map[string]float64, spans []correlation.SpanRef,
	signals []correlation.SignalRef, decompose bool) ([]EnrichmentResult, DebugStats) {
	in := gpu.NewInterner()
	sp := make([]gpu.Span, len(spans))
	ev := make([]gpu.Event, len(signals))
	base := make([]float64, len(spans)*gpu.NSig)
	for i, s := range spans {
		var err error
		if sp[i], err = in.Span(s.TraceID, s.Service, s.Node, s.Pod, s.PID, s.ConnTuple, s.Timestamp, gpu.None); err != nil {
			panic(err)
		}
		for k, key := range attrKeys {
			if v, ok := bases[i][key]; ok {
				base[i*gpu.NSig+k] = v
			} else {
				base[i*gpu.NSig+k] = math.NaN()
			}
		}
	}
	for i, g := range signals {
		var err error
		if ev[i], err = in.Event(g.Signal, g.TraceID, g.Service, g.Node, g.Pod, g.PID, g.ConnTuple, g.Timestamp, g.Value); err != nil {
			panic(err)
		}
	}
	w, thr, fan := c.params()
	r, err := gpu.MustDefault().ProcessBatch(sp, ev, in.Idents(), w, thr, fan, base)
	if err != nil {
		panic(err)
	}
	out := make([]EnrichmentResult, len(spans))
	for i := range spans {
		attrs := make(map[string]float64, len(bases[i])+4) // a fresh clone, correlator.go:69
		for k, v := range bases[i] {
			attrs[k] = v
		}
		res := EnrichmentResult{Attributes: attrs}
		for f := 0; f < r.Fanout; f++ {
			e := r.CandEvent[i*r.Fanout+f]
			if e == gpu.None {
				break
			}
			t := r.CandTier[i*r.Fanout+f]
			res.Candidates = append(res.Candidates, Candidate{Signal: signals[e],
				Decision: correlation.Decision{Matched: true, Confidence: gpu.TierConfidence[t], Tier: gpu.Tiers[t]}})
		}
		if len(res.Candidates) > 0 { // correlator.go:106-118
			for k, key := range attrKeys {
				if v := r.Attr[i*gpu.NSig+k]; !math.IsNaN(v) {
					attrs[key] = v
				}
			}
			attrs[semconv.AttrCorrelationConf] = r.Conf[i]
		}
		if decompose && r.Retrieval[i] > 0 { // DecomposeRetrieval, correlator.go:179-194
			attrs[semconv.AttrRetrievalKernelMS] = r.Retrieval[i]
		}
		out[i] = res
	}
	return out, DebugStats{Unmatched: int(r.Debug.Unmatched), LowConfidence: int(r.Debug.LowConfidence),
		FanoutDropped: int(r.Debug.FanoutDropped), UnsupportedType: int(r.Debug.UnsupportedType)}
}

// EnrichAttributes enriches a span from any supported signal set (GPU build of correlator.go:50-125).
func (c Correlator) EnrichAttributes(base map[string]float64, span correlation.SpanRef,
	signals []correlation.SignalRef) EnrichmentResult {
	if base == nil {
		base = map[string]float64{}
	}
	res, dbg := c.enrichBatch([]map[string]float64{base}, []correlation.SpanRef{span}, signals, false)
	res[0].Debug = dbg
	return res[0]
}

// ProcessBatch applies correlation enrichment over a span batch (GPU build of processor.go:29-49): one call for
// the whole batch instead of one EnrichAttributes per span.
func (c Correlator) ProcessBatch(spans []SpanRecord, signals []correlation.SignalRef) ProcessedBatch {
	refs := make([]correlation.SpanRef, len(spans))
	bases := make([]map[string]float64, len(spans))
	for i, item := range spans {
		refs[i] = correlation.SpanRef{TraceID: item.TraceID, Service: item.Service, Node: item.Node, Pod: item.Pod,
			PID: item.PID, ConnTuple: item.ConnTuple, Timestamp: item.Timestamp}
		bases[i] = item.Attributes
	}
	res, dbg := c.enrichBatch(bases, refs, signals, true)
	result := ProcessedBatch{Spans: make([]SpanRecord, 0, len(spans)), Debug: dbg}
	for i, item := range spans {
		item.Attributes = res[i].Attributes
		result.Spans = append(result.Spans, item)
	}
	return result
}
