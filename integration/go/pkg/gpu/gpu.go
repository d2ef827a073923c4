//go:build gpu

// Package gpu binds include/llmslo_gpu.h.  It knows nothing about the reference's own types (pkg/correlation,
// pkg/attribution import it, not the other way round).
package gpu

/*
#cgo LDFLAGS: -lllmslo_gpu
#include <stdlib.h>
#include "llmslo_gpu.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"sync"
	"unsafe"
)

// Memory layout mirrors of the C records (checked against lsg_abi_sizes in New).
type Event struct {
	TsNs     int64
	Value    float64
	Ident    uint32
	PID      uint32
	Conn     uint32
	TraceSig uint32 // trace id << 4 | signal index
}

type Span struct {
	TsNs     int64
	Ident    uint32
	PID      uint32
	Conn     uint32
	Trace    uint32
	Incident uint32
	Reserved uint32
}

type Ident struct{ Pod, Node, Service, Grp uint32 }

type Debug struct{ Unmatched, LowConfidence, FanoutDropped, UnsupportedType uint64 }

type Eval struct {
	TP, FP, FN, TN              uint64
	TierCorrect, TierComparable uint64
	PredictedByTier             [5]uint64 // predicted pairs per tier code
	MatchedByTier               [5]uint64 // all pairs per Match() tier code, [0] = unmatched
}

const (
	NSig           = int(C.LSG_NSIG)
	NDom           = int(C.LSG_NDOM)
	SigUnsupported = uint32(C.LSG_SIG_UNSUPPORTED)
	None           = uint32(0xffffffff)
	TraceBits      = 28
	MaxActual      = int(C.LSG_MAX_ACTUAL)
)

// TsZero marks time.Time{} (IsZero), which never matches (dns.go:107-113).
var TsZero = int64(C.LSG_TS_ZERO)

// Signals in DefaultLikelihoods order (bayesian.go:69-189); the index is the signal's slot everywhere in the ABI.
var Signals = [...]string{"dns_latency_ms", "tcp_retransmits_total", "runqueue_delay_ms", "connect_latency_ms",
	"tls_handshake_ms", "cpu_steal_pct", "cfs_throttled_ms", "mem_reclaim_latency_ms", "disk_io_latency_ms",
	"syscall_latency_ms", "connect_errors_total", "tls_handshake_fail_total"}

// Domains in AllDomains order (bayesian.go:23-34).
var Domains = [...]string{"network_dns", "network_egress", "cpu_throttle", "memory_pressure", "provider_throttle",
	"provider_error", "retrieval_backend", "unknown"}

// Tier names by tier code (dns.go:58-73); code 0 = no match.
var Tiers = [...]string{"", "trace_id_exact", "pod_pid_100ms", "pod_conn_250ms", "service_node_500ms"}

// TierConfidence by tier code (dns.go:58-73).
var TierConfidence = [...]float64{0, 1.0, 0.9, 0.8, 0.65}

// Ctx wraps lsg_ctx.  A context is single-threaded: calls run under the mutex on a locked OS thread so the CUDA
// primary context stays on one thread.  cgo calls block the goroutine, which is fine for batch work.
type Ctx struct {
	mu sync.Mutex
	h  *C.lsg_ctx
}

var (
	defaultOnce sync.Once
	defaultCtx  *Ctx
	defaultErr  error
)

// Default returns the process-wide context on device 0 (what the *_gpu.go files use).
func Default() (*Ctx, error) {
	defaultOnce.Do(func() { defaultCtx, defaultErr = New(0) })
	return defaultCtx, defaultErr
}

// MustDefault panics when there is no usable GPU: the gpu build has no CPU fallback.
func MustDefault() *Ctx {
	c, err := Default()
	if err != nil {
		panic(err)
	}
	return c
}

func New(device int) (*Ctx, error) {
	var sizes [12]C.size_t
	C.lsg_abi_sizes(&sizes[0], C.int(len(sizes)))
	if uintptr(sizes[0]) != unsafe.Sizeof(Event{}) || uintptr(sizes[1]) != unsafe.Sizeof(Span{}) ||
		uintptr(sizes[2]) != unsafe.Sizeof(Ident{}) || uintptr(sizes[3]) != unsafe.Sizeof(Debug{}) ||
		uintptr(sizes[4]) != unsafe.Sizeof(Eval{}) {
		return nil, fmt.Errorf("llmslo_gpu: struct layout differs from the library's (%s)", C.GoString(C.lsg_version()))
	}
	var h *C.lsg_ctx
	if rc := C.lsg_create(C.int(device), nil, &h); rc != 0 {
		return nil, fmt.Errorf("llmslo_gpu: lsg_create failed with code %d (no CPU fallback)", int(rc))
	}
	return &Ctx{h: h}, nil
}

func (c *Ctx) Close() {
	c.mu.Lock()
	defer c.mu.Unlock()
	if c.h != nil {
		C.lsg_destroy(c.h)
		c.h = nil
	}
}

func (c *Ctx) err(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return fmt.Errorf("llmslo_gpu: %s (code %d)", C.GoString(C.lsg_last_error(c.h)), int(rc))
}

func (c *Ctx) enter() func() {
	c.mu.Lock()
	runtime.LockOSThread()
	return func() { runtime.UnlockOSThread(); c.mu.Unlock() }
}

func ptr[T any](s []T) unsafe.Pointer {
	if len(s) == 0 {
		return nil
	}
	return unsafe.Pointer(unsafe.SliceData(s))
}

// MatchPairs is Match (dns.go:50-76) over pre-paired records plus the counters of EvaluateLabeledPairs
// (evaluator.go:90-167).  expectedTier: 0 = no expectation.
func (c *Ctx) MatchPairs(spans []Span, events []Event, idents []Ident, windowNS int64, threshold float64,
	expectedMatch, expectedTier []uint8) (tier, predicted []uint8, ev Eval, err error) {
	defer c.enter()()
	n := len(spans)
	tier, predicted = make([]uint8, n), make([]uint8, n)
	rc := C.lsg_match_pairs(c.h, (*C.lsg_span)(ptr(spans)), (*C.lsg_event)(ptr(events)), C.size_t(n),
		(*C.lsg_ident)(ptr(idents)), C.size_t(len(idents)), C.int64_t(windowNS), C.double(threshold),
		(*C.uint8_t)(ptr(expectedMatch)), (*C.uint8_t)(ptr(expectedTier)), (*C.uint8_t)(ptr(tier)),
		(*C.uint8_t)(ptr(predicted)), (*C.lsg_eval)(unsafe.Pointer(&ev)))
	return tier, predicted, ev, c.err(rc)
}

// BatchResult is what ProcessBatch needs to rebuild the attribute maps (processor.go:29-49).
type BatchResult struct {
	Fanout    int
	CandEvent []uint32  // n_spans * Fanout, None = empty slot; kept candidates in the reference's order
	CandTier  []uint8   // n_spans * Fanout
	Attr      []float64 // n_spans * NSig, NaN = absent; max(base, kept values) per signal (correlator.go:106-112)
	Conf      []float64 // llm.ebpf.correlation_confidence, 0 = not set (correlator.go:113-118)
	Retrieval []float64 // DecomposeRetrieval total (correlator.go:179-194), set when > 0
	Debug     Debug
}

// ProcessBatch is the body of Correlator.ProcessBatch.  attrBase (optional) holds the spans' existing values of
// the twelve eBPF attributes, NaN = absent.
func (c *Ctx) ProcessBatch(spans []Span, events []Event, idents []Ident, windowNS int64, threshold float64,
	fanout int, attrBase []float64) (BatchResult, error) {
	defer c.enter()()
	if fanout <= 0 {
		fanout = 3 // correlator.go:62-66
	}
	n := len(spans)
	r := BatchResult{Fanout: fanout, CandEvent: make([]uint32, n*fanout), CandTier: make([]uint8, n*fanout),
		Attr: make([]float64, n*NSig), Conf: make([]float64, n), Retrieval: make([]float64, n)}
	rc := C.lsg_process_batch(c.h, (*C.lsg_span)(ptr(spans)), C.size_t(n), (*C.lsg_event)(ptr(events)),
		C.size_t(len(events)), (*C.lsg_ident)(ptr(idents)), C.size_t(len(idents)), C.int64_t(windowNS),
		C.double(threshold), C.int(fanout), (*C.double)(ptr(attrBase)), (*C.uint32_t)(ptr(r.CandEvent)),
		(*C.uint8_t)(ptr(r.CandTier)), (*C.double)(ptr(r.Attr)), (*C.double)(ptr(r.Conf)),
		(*C.double)(ptr(r.Retrieval)), nil, (*C.lsg_debug)(unsafe.Pointer(&r.Debug)))
	return r, c.err(rc)
}

// Model is lsg_model: the flattened BayesianAttributor (bayesian.go:36-44).  Lik is NaN where a domain is missing
// from a signal's table (likelihoodFor then uses 0.5, bayesian.go:293-296).
type Model struct {
	Prior    [8]float64
	Lik      [12][8]float64
	Thresh   [12]float64
	LogExtra [8]float64
	SigMask  uint32
	Reserved uint32
}

func DefaultModel() Model {
	var m Model
	C.lsg_default_model((*C.lsg_model)(unsafe.Pointer(&m)))
	return m
}

func (c *Ctx) SetModel(m *Model) error {
	defer c.enter()()
	return c.err(C.lsg_set_model(c.h, (*C.lsg_model)(unsafe.Pointer(m))))
}

// Attribute is BayesianAttributor.Attribute for a batch: signals is n*NSig (NaN = absent), hasSignals marks
// samples with len(Signals) > 0 (the others keep ruleLabel, bayesian.go:318-320).  posterior is n*NDom in
// AllDomains order; top1 the first maximum (sort.SliceStable, bayesian.go:279-281); elevMask bit s = signal s
// elevated (bayesian.go:220-226).
func (c *Ctx) Attribute(signals []float64, hasSignals, ruleLabel []uint8) (posterior []float64, top1 []uint8,
	elevMask []uint16, err error) {
	defer c.enter()()
	n := len(signals) / NSig
	posterior, top1, elevMask = make([]float64, n*NDom), make([]uint8, n), make([]uint16, n)
	rc := C.lsg_attribute(c.h, (*C.double)(ptr(signals)), (*C.uint8_t)(ptr(hasSignals)), (*C.uint8_t)(ptr(ruleLabel)),
		C.size_t(n), (*C.double)(ptr(posterior)), (*C.uint8_t)(ptr(top1)), (*C.uint16_t)(ptr(elevMask)))
	return posterior, top1, elevMask, c.err(rc)
}

// Confusion is lsg_confusion_out.
type Confusion struct {
	Matrix   [16][8]uint64
	N        uint64
	Correct  uint64
	Partial  uint64
	Coverage [9][9]uint64 // [found][len(expected)]
}

// ConfusionCounts is BuildConfusionMatrix + the integer numerators of Accuracy / PartialAccuracy /
// CoverageAccuracy (pipeline.go:65-185).  actual: 0..7 = AllDomains, 8..15 = other strings (numbered by the caller);
// expected: up to eight 4-bit domain indices per incident, first in the low nibble, 0xF = not one of AllDomains;
// expectedN: how many; posterior: n*NDom, NaN where the prediction carries no hypothesis for the domain.
func (c *Ctx) ConfusionCounts(actual, predicted []uint8, expected []uint32, expectedN []uint8, posterior []float64,
	coverageThr, hypFloor float64) (Confusion, error) {
	defer c.enter()()
	var out Confusion
	rc := C.lsg_confusion(c.h, (*C.uint8_t)(ptr(actual)), (*C.uint8_t)(ptr(predicted)), (*C.uint32_t)(ptr(expected)),
		(*C.uint8_t)(ptr(expectedN)), (*C.double)(ptr(posterior)), C.double(coverageThr), C.double(hypFloor),
		C.size_t(len(actual)), (*C.lsg_confusion_out)(unsafe.Pointer(&out)))
	return out, c.err(rc)
}

// SegmentQuantiles is quantile() (pkg/slo/calculator.go:121-149) for many value lists in one call.
func (c *Ctx) SegmentQuantiles(values []float64, segment []uint32, nSegments int, q []float64) ([]float64, []uint64, error) {
	defer c.enter()()
	out, cnt := make([]float64, nSegments*len(q)), make([]uint64, nSegments)
	rc := C.lsg_segment_quantiles(c.h, (*C.double)(ptr(values)), (*C.uint32_t)(ptr(segment)), C.size_t(len(values)),
		C.size_t(nSegments), (*C.double)(ptr(q)), C.size_t(len(q)), (*C.double)(ptr(out)), (*C.uint64_t)(ptr(cnt)))
	return out, cnt, c.err(rc)
}

// PipelineResult is what one lsg_pipeline call returns for a batch of spans grouped into incidents: the kept candidates
// per span in compact form (event index, tier, signal, value: a span's attribute map is max(value) per signal over
// them, its confidence TierConfidence[CandTier[first]]), the per-incident signal maxima, posteriors and top-1 labels,
// and the batch counters (debug stats, confusion matrix, accuracy numerators).
type PipelineResult struct {
	Fanout     int
	CandEvent  []uint32
	CandTier   []uint8
	CandSig    []uint8 // 0xff = empty slot
	CandValue  []float64
	IncSignals []float64 // n_incidents * NSig, NaN = absent
	Posterior  []float64 // n_incidents * NDom
	Top1       []uint8
	Counters   Counters
}

// Counters is lsg_counters.
type Counters struct {
	Debug                                   Debug
	Confusion                               Confusion
	Events, Spans, Incidents, Candidates, Kept uint64
	BadIdents                               uint64
	UnsupportedEvents                       uint64
}

// KeyHint is the optional statement about a batch (lsg_pipeline_args.key_hint): every span that can match has its time in
// [SpanTsLo, SpanTsHi] and no identity row's group exceeds MaxGrp.  With it the library never synchronises mid-call.
type KeyHint struct {
	SpanTsLo, SpanTsHi int64
	MaxGrp             uint32
}

// Pipeline runs ProcessBatch + the per-incident maxima + BuildAttributionsBayes + BuildConfusionMatrix for one batch in
// one call (processor.go:29, pipeline.go:31,65).  Spans carry their incident index; incActual / incExpected /
// incExpectedN are encoded as for ConfusionCounts.
func (c *Ctx) Pipeline(spans []Span, events []Event, idents []Ident, incActual []uint8, incExpected []uint32,
	incExpectedN []uint8, windowNS int64, threshold float64, fanout int, coverageThr float64, hint *KeyHint) (PipelineResult, error) {
	defer c.enter()()
	if fanout <= 0 {
		fanout = 3
	}
	ns, ni := len(spans), len(incActual)
	r := PipelineResult{Fanout: fanout, CandEvent: make([]uint32, ns*fanout), CandTier: make([]uint8, ns*fanout),
		CandSig: make([]uint8, ns*fanout), CandValue: make([]float64, ns*fanout), IncSignals: make([]float64, ni*NSig),
		Posterior: make([]float64, ni*NDom), Top1: make([]uint8, ni)}
	// `a` holds pointers into Go slices, so passing &a alone would break the cgo pointer rule ("Go pointer to
	// unpinned Go pointer"): every slice the struct names is pinned for the duration of the call (Go >= 1.21).
	var pin runtime.Pinner
	defer pin.Unpin()
	pinned := func(p unsafe.Pointer) unsafe.Pointer {
		if p != nil {
			pin.Pin(p)
		}
		return p
	}
	for _, p := range []unsafe.Pointer{ptr(spans), ptr(events), ptr(idents), ptr(incActual), ptr(incExpected),
		ptr(incExpectedN), ptr(r.CandEvent), ptr(r.CandTier), ptr(r.CandSig), ptr(r.CandValue), ptr(r.IncSignals),
		ptr(r.Posterior), ptr(r.Top1)} {
		pinned(p)
	}
	var a C.lsg_pipeline_args
	a.spans, a.n_spans = (*C.lsg_span)(ptr(spans)), C.size_t(ns)
	a.events, a.n_events = (*C.lsg_event)(ptr(events)), C.size_t(len(events))
	a.idents, a.n_idents = (*C.lsg_ident)(ptr(idents)), C.size_t(len(idents))
	a.n_incidents = C.size_t(ni)
	a.inc_actual, a.inc_expected = (*C.uint8_t)(ptr(incActual)), (*C.uint32_t)(ptr(incExpected))
	a.inc_expected_n = (*C.uint8_t)(ptr(incExpectedN))
	a.window_ns, a.threshold, a.fanout, a.coverage_thr = C.int64_t(windowNS), C.double(threshold), C.int(fanout), C.double(coverageThr)
	a.cand_event, a.cand_tier = (*C.uint32_t)(ptr(r.CandEvent)), (*C.uint8_t)(ptr(r.CandTier))
	a.cand_sig, a.cand_value = (*C.uint8_t)(ptr(r.CandSig)), (*C.double)(ptr(r.CandValue))
	a.inc_signals, a.posterior, a.top1 = (*C.double)(ptr(r.IncSignals)), (*C.double)(ptr(r.Posterior)), (*C.uint8_t)(ptr(r.Top1))
	if hint != nil {
		a.key_hint, a.max_grp_hint = 1, C.uint32_t(hint.MaxGrp)
		a.span_ts_lo, a.span_ts_hi = C.int64_t(hint.SpanTsLo), C.int64_t(hint.SpanTsHi)
	}
	rc := C.lsg_pipeline(c.h, &a, (*C.lsg_counters)(unsafe.Pointer(&r.Counters)))
	return r, c.err(rc)
}

// ---- round 2 additions to the C ABI -------------------------------------------------------------------------------

// Event24 is lsg_event24: the 24-byte upload form of an Event (time offset from a per-batch base, narrow ids).
type Event24 struct {
	Value float64
	A, B  uint64
}

// PackEvents24 packs 32-byte records for the upload (lsg_pack_events24).  misfit > 0: some record does not fit the narrow
// fields and the batch has to go up as 32-byte records.  An interner can emit Event24 directly and skip this pass.
func PackEvents24(events []Event, base int64) (out []Event24, misfit int64) {
	out = make([]Event24, len(events))
	misfit = int64(C.lsg_pack_events24((*C.lsg_event)(ptr(events)), C.size_t(len(events)), C.int64_t(base),
		(*C.lsg_event24)(ptr(out))))
	return out, misfit
}

// SetJoinPath selects the implementation of the sort + join stages (0 permutation sort + gathering join, 1 bucket
// partition + slab join, -1 the LSG_JOIN_PATH environment variable decides; 2 / 3 are path 0 with the sliding-tile /
// the fixed-chunk form of its tile kernel forced instead of chosen by the batch's events per span).
func (c *Ctx) SetJoinPath(path int) error {
	defer c.enter()()
	return c.err(C.lsg_set_join_path(c.h, C.int(path)))
}

// ValidateGroups checks the join-group contract on a batch (lsg_validate_groups): identity rows whose pod /
// (service, node) pair, and records whose trace id, were first seen under another group; idents outside the table.
// All zero = the contract holds.  The interner's own output always passes; use it on batches packed elsewhere.
func (c *Ctx) ValidateGroups(spans []Span, events []Event, idents []Ident) (pod, class, trace, badIdent uint64, err error) {
	defer c.enter()()
	var out [4]C.uint64_t
	rc := C.lsg_validate_groups(c.h, (*C.lsg_span)(ptr(spans)), C.size_t(len(spans)), (*C.lsg_event)(ptr(events)),
		C.size_t(len(events)), (*C.lsg_ident)(ptr(idents)), C.size_t(len(idents)), &out[0])
	return uint64(out[0]), uint64(out[1]), uint64(out[2]), uint64(out[3]), c.err(rc)
}

// Multi wraps lsg_multi: one context per GPU of the box, NCCL allreduce of the counter block inside the library.
type Multi struct {
	mu sync.Mutex
	h  *C.lsg_multi
	n  int
}

// NewMulti creates one context per listed device (distinct ordinals).
func NewMulti(devices []int) (*Multi, error) {
	dv := make([]C.int, len(devices))
	for i, d := range devices {
		dv[i] = C.int(d)
	}
	var h *C.lsg_multi
	if rc := C.lsg_multi_create(&dv[0], C.int(len(dv)), &h); rc != 0 {
		msg := "invalid device list"
		if h != nil {
			msg = C.GoString(C.lsg_multi_last_error(h))
			C.lsg_multi_destroy(h)
		}
		return nil, fmt.Errorf("llmslo_gpu: %s (code %d)", msg, int(rc))
	}
	return &Multi{h: h, n: len(devices)}, nil
}

func (m *Multi) Close() {
	m.mu.Lock()
	defer m.mu.Unlock()
	if m.h != nil {
		C.lsg_multi_destroy(m.h)
		m.h = nil
	}
}

// Shard is one rank's part of a batch: no join group may be split between shards (shard by node / pod identity).
type Shard struct {
	Spans                     []Span
	Events                    []Event
	Idents                    []Ident
	IncActual, IncExpectedN   []uint8
	IncExpected               []uint32
}

// Pipeline runs the whole path on every shard at once and returns each rank's results and the batch's counters
// (lsg_multi_pipeline: the sum of the ranks' counters with Unmatched / UnsupportedType on the whole batch's footing,
// i.e. what ProcessBatch + BuildConfusionMatrix return for the concatenated batch).
func (m *Multi) Pipeline(shards []Shard, windowNS int64, threshold float64, fanout int, coverageThr float64) ([]PipelineResult, Counters, error) {
	m.mu.Lock()
	runtime.LockOSThread()
	defer func() { runtime.UnlockOSThread(); m.mu.Unlock() }()
	if len(shards) != m.n {
		return nil, Counters{}, fmt.Errorf("llmslo_gpu: %d shards for %d devices", len(shards), m.n)
	}
	if fanout <= 0 {
		fanout = 3
	}
	var pin runtime.Pinner
	defer pin.Unpin()
	keep := func(p unsafe.Pointer) unsafe.Pointer {
		if p != nil {
			pin.Pin(p)
		}
		return p
	}
	// the argument structs live in C memory, so only pinned Go memory is referenced from them
	args := (*[1 << 10]C.lsg_pipeline_args)(C.calloc(C.size_t(m.n), C.size_t(unsafe.Sizeof(C.lsg_pipeline_args{}))))
	defer C.free(unsafe.Pointer(args))
	res := make([]PipelineResult, m.n)
	for r, sh := range shards {
		ns, ni := len(sh.Spans), len(sh.IncActual)
		res[r] = PipelineResult{Fanout: fanout, CandEvent: make([]uint32, ns*fanout), CandTier: make([]uint8, ns*fanout),
			CandSig: make([]uint8, ns*fanout), CandValue: make([]float64, ns*fanout), IncSignals: make([]float64, ni*NSig),
			Posterior: make([]float64, ni*NDom), Top1: make([]uint8, ni)}
		a := &args[r]
		a.spans, a.n_spans = (*C.lsg_span)(keep(ptr(sh.Spans))), C.size_t(ns)
		a.events, a.n_events = (*C.lsg_event)(keep(ptr(sh.Events))), C.size_t(len(sh.Events))
		a.idents, a.n_idents = (*C.lsg_ident)(keep(ptr(sh.Idents))), C.size_t(len(sh.Idents))
		a.n_incidents = C.size_t(ni)
		a.inc_actual, a.inc_expected = (*C.uint8_t)(keep(ptr(sh.IncActual))), (*C.uint32_t)(keep(ptr(sh.IncExpected)))
		a.inc_expected_n = (*C.uint8_t)(keep(ptr(sh.IncExpectedN)))
		a.window_ns, a.threshold, a.fanout, a.coverage_thr = C.int64_t(windowNS), C.double(threshold), C.int(fanout), C.double(coverageThr)
		a.cand_event, a.cand_tier = (*C.uint32_t)(keep(ptr(res[r].CandEvent))), (*C.uint8_t)(keep(ptr(res[r].CandTier)))
		a.cand_sig, a.cand_value = (*C.uint8_t)(keep(ptr(res[r].CandSig))), (*C.double)(keep(ptr(res[r].CandValue)))
		a.inc_signals, a.posterior = (*C.double)(keep(ptr(res[r].IncSignals))), (*C.double)(keep(ptr(res[r].Posterior)))
		a.top1 = (*C.uint8_t)(keep(ptr(res[r].Top1)))
	}
	own := make([]Counters, m.n)
	var red Counters
	rc := C.lsg_multi_pipeline(m.h, &args[0], (*C.lsg_counters)(unsafe.Pointer(&own[0])), (*C.lsg_counters)(unsafe.Pointer(&red)))
	if rc != 0 {
		return nil, red, fmt.Errorf("llmslo_gpu: %s (code %d)", C.GoString(C.lsg_multi_last_error(m.h)), int(rc))
	}
	for r := range res {
		res[r].Counters = own[r]
	}
	return res, red, nil
}
