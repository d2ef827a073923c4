/*
 * llmslo_gpu.h — C ABI of libllmslo_gpu.so: the B200 (sm_100a) batch correlation + attribution path
 * of llm-slo-ebpf-toolkit.
 *
 * The reference (Go, CGO_ENABLED=0) has no FFI seam; the drop-in boundary is a set of exported Go
 * functions.  Each entry point below names the Go function(s) whose arithmetic it
 * replaces, as file:line under the reference tree.  The cgo binding a maintainer would add
 * (pkg/gpu) is shown in INTEGRATION.md.
 *
 * Conventions
 *   - plain pointers and sizes; no C++ types, no exceptions, no abort().  Return 0 on success, a
 *     negative LSG_ERR_* otherwise; lsg_last_error(ctx) gives the text until the next call on ctx.
 *   - the caller owns every buffer.  Functions without a suffix take HOST pointers and copy in/out
 *     on the context's stream (pinned host memory makes those copies asynchronous); `_dev`
 *     functions take DEVICE pointers on the context's device and leave results on the device.
 *     Every call returns after its work on the context's stream has completed, except `_dev`
 *     calls documented as asynchronous.
 *   - device record arrays (lsg_event, lsg_span) passed to `_dev` functions must be 32-byte aligned (they are
 *     read with 256-bit loads); cudaMalloc and framework allocators give at least that.
 *   - a context is bound to one device and one stream and is NOT thread-safe.
 *   - there is no CPU fallback: without a usable CUDA device lsg_create fails.
 *
 * Interned identities.  Strings are dictionary-encoded by the caller (0 = empty string).  A record
 * does not carry pod/node/service directly but an `ident`: a row of the identity table holding the
 * interned (pod, node, service) triple seen on that record, plus `grp`, the join group.
 *
 *   JOIN-GROUP CONTRACT (lsg_validate_groups checks it for a batch): any span/event pair for which the
 *   reference Match() (pkg/correlation/dns.go:50-76) can return Matched must have the same
 *   non-zero grp.  The interner gets this by union-find over identity rows: rows sharing a
 *   non-empty pod, rows sharing a non-empty (service, node), and rows that carry the same
 *   non-empty trace id are merged.  grp 0 means "matches nothing".  Within a group the library
 *   evaluates the complete four-tier Match() on every pair inside the outer window, so coarser
 *   groups are always correct, only slower.
 */
#ifndef LLMSLO_GPU_H
#define LLMSLO_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LSG_NSIG 12            /* supported signals, in DefaultLikelihoods order (bayesian.go:69-189)      */
#define LSG_NDOM 8             /* fault domains, AllDomains() order (bayesian.go:23-34)                    */
#define LSG_SIG_UNSUPPORTED 15 /* signalAttrKey() default branch (ebpfcorrelator/correlator.go:143-172)    */
#define LSG_TRACE_BITS 28      /* events pack trace id and signal into one word                            */
#define LSG_TS_ZERO INT64_MIN  /* time.Time{}.IsZero() (dns.go:108); Unix 0 is a valid time                */
#define LSG_NONE 0xFFFFFFFFu   /* "no event" / "no incident"                                               */
#define LSG_MAX_FANOUT 32     /* MaxJoinFanout accepted; up to 8 runs on the tile kernel, 9..32 on the warp-per-span kernel */
#define LSG_MAX_ACTUAL 16      /* rows of the confusion matrix: 8 domains + up to 8 other actual strings   */
#define LSG_MAX_WINDOW_NS (((int64_t)1 << 53) - 1) /* float64(Duration) compare in dns.go:111-112 is exact below this */

/* tiers of Match(), dns.go:58-73.  confidence = {0, 1.0, 0.9, 0.8, 0.65}[tier] */
enum { LSG_TIER_NONE = 0, LSG_TIER_TRACE = 1, LSG_TIER_POD_PID = 2, LSG_TIER_POD_CONN = 3, LSG_TIER_SVC_NODE = 4 };

enum {
  LSG_OK = 0,
  LSG_ERR_INVALID = -1,   /* bad argument (null pointer, fanout out of range, window too large, ...)  */
  LSG_ERR_CUDA = -2,      /* a CUDA runtime call failed; text in lsg_last_error                      */
  LSG_ERR_NOMEM = -3,     /* device or host allocation failed                                        */
  LSG_ERR_RANGE = -4,     /* an id in a record is outside its table (ident, trace >= 2^28, class ...) */
  LSG_ERR_NODEVICE = -5   /* no CUDA device: there is no CPU path                                     */
};

/* SignalRef (dns.go:30-40) after interning: 32 bytes. */
typedef struct {
  int64_t ts_ns;      /* Unix ns; LSG_TS_ZERO for the zero Time                                    */
  double value;       /* SignalRef.Value                                                          */
  uint32_t ident;     /* row of the identity table                                                */
  uint32_t pid;       /* PID if > 0 else 0                                                        */
  uint32_t conn;      /* interned ConnTuple                                                       */
  uint32_t trace_sig; /* (interned TraceID << 4) | signal index (0..11, or LSG_SIG_UNSUPPORTED)   */
} lsg_event;

/* The same record in 24 bytes, for the upload: end to end a batch costs its PCIe transfer, not its kernels, so the
 * form that crosses the bus is narrower than the one the kernels read (the first device pass widens it).  No reference
 * counterpart (SignalRef is a Go struct of strings); the interner emits one or the other.  Limits of the narrow fields:
 *   time      LSG_E24_TS_BITS-bit offset in ns from a per-batch `base` (73 minutes); all ones = the zero Time.  The
 *             caller chooses base <= first span - window and needs last span + window < base + 2^42 - 1 (checked: the
 *             pipeline refuses the batch otherwise); an event outside that range can match nothing and is encoded as
 *             the zero Time
 *   ident     < 2^14 identity rows,  pid < 2^22 (Linux PID_MAX_LIMIT),  trace id < 2^26,  conn id < 2^20 (ids are
 *             per batch, so these bound the distinct values of one batch)
 * lsg_pack_events24 reports how many records do not fit; a batch with any such record must go up in the 32-byte form. */
#define LSG_E24_TS_BITS 42
#define LSG_E24_IDENT_BITS 14
#define LSG_E24_PID_BITS 22
#define LSG_E24_TRACE_BITS 26
#define LSG_E24_CONN_BITS 20
typedef struct {
  double value;
  uint64_t a; /* time offset [0,42) | ident [42,56) | signal [56,60) | pid bits 0..3 [60,64) */
  uint64_t b; /* trace id [0,26) | conn [26,46) | pid bits 4..21 [46,64)                      */
} lsg_event24;

/* SpanRef (dns.go:19-27) / SpanRecord (processor.go:10-20) after interning: 32 bytes. */
typedef struct {
  int64_t ts_ns;
  uint32_t ident;
  uint32_t pid;
  uint32_t conn;
  uint32_t trace;     /* interned TraceID, < 2^28                                                  */
  uint32_t incident;  /* incident this span belongs to (A7 glue), or LSG_NONE                      */
  uint32_t reserved;  /* must be 0                                                                 */
} lsg_span;

/* identity table row */
typedef struct {
  uint32_t pod, node, service; /* interned strings, 0 = empty                                      */
  uint32_t grp;                /* join group, see contract above; 0 = matches nothing              */
} lsg_ident;

/* DebugStats (correlator.go:20-25), summed over spans as mergeDebug does (processor.go:51-58). */
typedef struct {
  uint64_t unmatched, low_confidence, fanout_dropped, unsupported_type;
} lsg_debug;

/* integer part of EvalReport (evaluator.go:34-51); ratios are formed by the caller. */
typedef struct {
  uint64_t tp, fp, fn, tn;
  uint64_t tier_correct, tier_comparable;
  uint64_t predicted_by_tier[5]; /* predicted pairs per tier; mean confidence = sum(n[t]*conf[t]) / sum(n[t]) */
  uint64_t matched_by_tier[5];   /* all pairs per Match() tier, [0] = unmatched                              */
} lsg_eval;

/* BayesianAttributor (bayesian.go:39-42) flattened.  NaN in lik = domain missing from that signal's
 * map (likelihoodFor returns 0.5, bayesian.go:293-296); NaN in thresh = no threshold. */
typedef struct {
  double prior[LSG_NDOM];          /* b.Priors in AllDomains order; <= 0 becomes 1e-10 (bayesian.go:231-233) */
  double lik[LSG_NSIG][LSG_NDOM];  /* b.Likelihoods[signal][domain]                                          */
  double thresh[LSG_NSIG];         /* signalThresholds (bayesian.go:193-206)                                 */
  double log_extra[LSG_NDOM];      /* sum of log clamp(1-p) over Likelihoods entries outside the 12 signals  */
  uint32_t sig_mask;               /* bit s set = signal s is a key of b.Likelihoods                         */
  uint32_t reserved;
} lsg_model;

/* integer part of Accuracy / PartialAccuracy / CoverageAccuracy and BuildConfusionMatrix (pipeline.go:65-185) */
typedef struct {
  uint64_t matrix[LSG_MAX_ACTUAL][LSG_NDOM]; /* [actual class][predicted domain]                   */
  uint64_t n;                                /* incidents counted                                  */
  uint64_t correct;                          /* Accuracy numerator                                 */
  uint64_t partial;                          /* PartialAccuracy numerator                          */
  uint64_t coverage[LSG_NDOM + 1][LSG_NDOM + 1]; /* [found][len(expected)] histogram; coverage accuracy = sum h*found/len / n */
} lsg_confusion_out;

/* per (incident, signal) features of the A7 glue.  Only `max` has a reference counterpart
 * (correlator.go:109-111); count/sum/sketch are defined by this library (and restated for the tests in oracle/oracle.cpp). */
#define LSG_SKETCH_BINS 16
typedef struct {
  uint32_t count;                    /* spans of the incident that carry the signal        */
  uint32_t sketch[LSG_SKETCH_BINS];  /* log2 histogram of the per-span values              */
  double sum;                        /* sum of the per-span values (span order)            */
} lsg_feature;

typedef struct lsg_ctx lsg_ctx;

/* ---- context ------------------------------------------------------------------------------ */
/* device: CUDA ordinal.  stream: a cudaStream_t (as void*) to run on, or NULL for a private one. */
int lsg_create(int device, void *stream, lsg_ctx **out);
void lsg_destroy(lsg_ctx *ctx);
const char *lsg_last_error(const lsg_ctx *ctx);
const char *lsg_version(void);
/* kernels launched by this context since creation (bench.py's gpu_launches) */
uint64_t lsg_launch_count(const lsg_ctx *ctx);
/* Which of the two implementations of the sort + join stages lsg_process_batch / lsg_pipeline take (same results):
 * 0 = permutation sort (counting / radix) + a join that gathers its tiles through the permutation (default),
 * 1 = bucket partition of the records + slab join on contiguous tiles, -1 = the LSG_JOIN_PATH environment variable
 * ("slab" or "perm") decides.  Path 0 picks one of two forms of its tile kernel by the batch's density (a sliding tile
 * when a chunk of spans reaches more events than one tile holds, fixed chunks otherwise); 2 and 3 are path 0 with the
 * sliding / the fixed form forced (tests, measurements).  A tuning knob with no reference counterpart. */
int lsg_set_join_path(lsg_ctx *ctx, int path);
/* bytes of device workspace currently held */
size_t lsg_workspace_bytes(const lsg_ctx *ctx);

/* Page-locked host memory for the buffers handed to the host entry points (lsg_pipeline, lsg_pipeline_submit,
 * lsg_process_batch ...): copies from pageable memory are staged by the driver and cannot overlap.  write_combined
 * != 0 asks for write-combining memory (cudaHostAllocWriteCombined): faster for the device to read on some hosts,
 * very slow for the CPU to read back, so only for input buffers the host fills once and never reads.  Free with
 * lsg_host_free.  No reference counterpart (Go slices are pageable); pkg/gpu would keep these as its staging. */
int lsg_host_alloc(size_t bytes, int write_combined, void **out);
void lsg_host_free(void *p);

/* sizeof() of the ABI structs as compiled into the library, in header order (event, span, ident, debug, eval,
 * model, confusion_out, feature, counters, pipeline_args, synth_cfg, event24); returns how many there are.  Lets a
 * binding check its own struct mirrors at load time. */
int lsg_abi_sizes(size_t *out, int n);

/* NewBayesianAttributor() defaults (bayesian.go:46-207) */
void lsg_default_model(lsg_model *out);
/* replaces reading b.Priors / b.Likelihoods inside Attribute (bayesian.go:230-242) */
int lsg_set_model(lsg_ctx *ctx, const lsg_model *model);

/* ---- A2-A3: correlation.Match (dns.go:50-76) over pre-paired records, and
 *      correlation.EvaluateLabeledPairs (evaluator.go:90-167) ------------------------------- */
int lsg_match_pairs(lsg_ctx *ctx, const lsg_span *spans, const lsg_event *events, size_t n,
                    const lsg_ident *idents, size_t n_idents, int64_t window_ns, double threshold,
                    const uint8_t *expected_match /* n or NULL */, const uint8_t *expected_tier /* n or NULL; 0 = none */,
                    uint8_t *tier_out /* n */, uint8_t *predicted_out /* n */, lsg_eval *eval_out);
int lsg_match_pairs_dev(lsg_ctx *ctx, const lsg_span *spans, const lsg_event *events, size_t n,
                        const lsg_ident *idents, size_t n_idents, int64_t window_ns, double threshold,
                        const uint8_t *expected_match, const uint8_t *expected_tier,
                        uint8_t *tier_out, uint8_t *predicted_out, lsg_eval *eval_out /* HOST */);

/* ---- A4-A6: Correlator.ProcessBatch (ebpfcorrelator/processor.go:29-49) =
 *      EnrichAttributes (correlator.go:50-125) per span + DecomposeRetrieval (correlator.go:179-194)
 *
 *  window_ns <= 0 -> 2 s; threshold <= 0 -> 0.7; fanout <= 0 -> 3 (dns.go:51-53, correlator.go:58-66).
 *  cand_event  [n_spans*fanout] index into `events` of the kept candidates in reference order
 *              (confidence desc, |dt| asc, input order), LSG_NONE = none
 *  cand_tier   [n_spans*fanout] tier of each kept candidate, 0 = none
 *  attr_base   [n_spans*12] or NULL: SpanRecord.Attributes on entry, NaN = absent.  NaN is this ABI's "absent": an
 *              event VALUE that is NaN is outside the parity claim (the reference keeps an existing NaN,
 *              correlator.go:109; here a later candidate replaces it)
 *  attr_out    [n_spans*12] enriched attributes, NaN = absent (correlator.go:106-112)
 *  conf_out    [n_spans] llm.ebpf.correlation_confidence, 0 = not set (correlator.go:113-118)
 *  retrieval_out [n_spans] or NULL: DecomposeRetrieval total (attribute is set when > 0)
 *  n_cand_out  [n_spans] or NULL: candidates >= threshold before the fanout cut
 *  Any output pointer except cand_event/cand_tier/attr_out/conf_out may be NULL. */
int lsg_process_batch(lsg_ctx *ctx, const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                      const lsg_ident *idents, size_t n_idents, int64_t window_ns, double threshold, int fanout,
                      const double *attr_base, uint32_t *cand_event, uint8_t *cand_tier, double *attr_out,
                      double *conf_out, double *retrieval_out, uint32_t *n_cand_out, lsg_debug *debug_out);
int lsg_process_batch_dev(lsg_ctx *ctx, const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                          const lsg_ident *idents, size_t n_idents, int64_t window_ns, double threshold, int fanout,
                          const double *attr_base, uint32_t *cand_event, uint8_t *cand_tier, double *attr_out,
                          double *conf_out, double *retrieval_out, uint32_t *n_cand_out, lsg_debug *debug_out /* HOST */);

/* the sort stage alone (device pointers): stable sort of events by (grp, ts); writes the sorted
 * records and, if not NULL, the permutation (sorted position -> input index) and the sort keys. */
int lsg_sort_events_dev(lsg_ctx *ctx, const lsg_event *events, size_t n_events, const lsg_ident *idents, size_t n_idents,
                        lsg_event *sorted_out, uint32_t *perm_out, uint64_t *keys_out);

/* ---- A7 glue: per incident, per signal max over its spans' enriched attributes
 *      (aggregate = max as in correlator.go:109-111; FaultSample.Signals, mapper.go:20).
 *  span_incident [n_spans] incident index or LSG_NONE;  signals_out [n_incidents*12], NaN = absent;
 *  features_out [n_incidents*12] or NULL. */
int lsg_incident_features(lsg_ctx *ctx, const double *span_attr, const uint32_t *span_incident, size_t n_spans,
                          size_t n_incidents, double *signals_out, lsg_feature *features_out);
int lsg_incident_features_dev(lsg_ctx *ctx, const double *span_attr, const uint32_t *span_incident, size_t n_spans,
                              size_t n_incidents, double *signals_out, lsg_feature *features_out);

/* No counterpart in the reference.  The same per-incident maxima as lsg_incident_features_dev (signals only, no
 * count/sum/sketch), for callers that must not allocate device memory: a process whose whole GPU memory budget
 * already belongs to its own pool, or a service that cannot take the device-wide synchronisation cudaMalloc and
 * cudaFree imply while other streams are running.  It needs no lsg_ctx and calls no cudaMalloc, cudaFree or other
 * allocating routine: the accumulators are a fixed array inside the library image (LSG_STATIC_MAX_INCIDENTS x 12
 * x 8 bytes = 1.5 MiB of device memory per GPU, taken when the driver first loads this part of the library on that
 * GPU - under CUDA's default lazy loading that is this function's first call there - and held until the process
 * exits).  Make that first call on each device before handing the remaining device memory to your own pool
 * (or run with CUDA_MODULE_LOADING=EAGER).
 *   - all pointers are DEVICE pointers on the CURRENT device; `stream` is a cudaStream_t of that device (NULL = the
 *     default stream).  The work is enqueued on `stream` and the call returns without synchronising.
 *   - n_incidents > LSG_STATIC_MAX_INCIDENTS is refused with LSG_ERR_RANGE; split the batch or use the
 *     allocating entry point.
 *   - the first call on each device looks up the array's address on that device; make that first call before
 *     starting any stream capture.  Device ordinals 0..63 are supported (LSG_ERR_RANGE beyond).
 *   - the return value also reports a sticky CUDA error left on the device by an earlier, unrelated launch.
 *   - every call on a device shares the one array and begins by clearing the part it uses, so the previous call's
 *     values never leak in, but two calls whose work overlaps on the same GPU (different streams or threads)
 *     race on it: keep calls on one device stream-ordered.
 *   - errbuf (optional, errbuf_len bytes) receives a message when the return value is not 0. */
#define LSG_STATIC_MAX_INCIDENTS 16384
int lsg_incident_signals_static(void *stream, const double *span_attr, const uint32_t *span_incident, size_t n_spans,
                                size_t n_incidents, double *signals_out, char *errbuf, size_t errbuf_len);

/* ---- A9-A10: BayesianAttributor.Attribute (bayesian.go:218-284) and the numeric part of
 *      AttributeSample (bayesian.go:315-343) for a batch of incidents.
 *  signals     [n*12] FaultSample.Signals by signal index, NaN = key absent
 *  has_signals [n] or NULL: len(sample.Signals) > 0 (NULL = derive: any non-NaN)
 *  rule_label  [n] or NULL: MapFaultLabel(sample.FaultLabel) as a domain index, used when !has_signals
 *  posterior   [n*8] AllDomains order (all NaN when !has_signals)
 *  top1        [n] predicted domain: first maximum of the posteriors in AllDomains order
 *              (sort.SliceStable, bayesian.go:279-281), or rule_label (0xFF if NULL) when !has_signals
 *  elev_mask   [n] or NULL: bit s = signal s elevated (bayesian.go:220-225) */
int lsg_attribute(lsg_ctx *ctx, const double *signals, const uint8_t *has_signals, const uint8_t *rule_label, size_t n,
                  double *posterior, uint8_t *top1, uint16_t *elev_mask);
int lsg_attribute_dev(lsg_ctx *ctx, const double *signals, const uint8_t *has_signals, const uint8_t *rule_label, size_t n,
                      double *posterior, uint8_t *top1, uint16_t *elev_mask);

/* ---- A12-A13: BuildConfusionMatrix, Accuracy, PartialAccuracy, CoverageAccuracy (pipeline.go:65-185)
 *  actual      [n] class of ExpectedDomain or MapFaultLabel(FaultLabel): 0..7 = AllDomains, 8..15 = other strings
 *  predicted   [n] PredictedFaultDomain, 0..7
 *  expected    [n] ExpectedDomains as nibbles (first element in the low nibble; 0..7 domain, 0xF other), or
 *              the one-element fallback list of pipeline.go:119-126
 *  expected_n  [n] length of that list, 1..8
 *  posterior   [n*8] or NULL (no hypotheses, the rule path); entries NaN = no hypothesis
 *  coverage_thr threshold of CoverageAccuracy (pipeline.go:163-167)
 *  hyp_floor   entries below it are not hypotheses: 0.01 when `posterior` is the raw output of lsg_attribute
 *              (AttributeSample's filter, bayesian.go:326-328), 0 when it already is a hypothesis list */
int lsg_confusion(lsg_ctx *ctx, const uint8_t *actual, const uint8_t *predicted, const uint32_t *expected,
                  const uint8_t *expected_n, const double *posterior, double coverage_thr, double hyp_floor,
                  size_t n, lsg_confusion_out *out);
int lsg_confusion_dev(lsg_ctx *ctx, const uint8_t *actual, const uint8_t *predicted, const uint32_t *expected,
                      const uint8_t *expected_n, const double *posterior, double coverage_thr, double hyp_floor,
                      size_t n, lsg_confusion_out *out /* HOST */);

/* ---- the whole hot path on device-resident input (bench.py's timed step):
 *      sort -> join -> per-span features -> per-incident features -> attribution -> confusion.
 *  `counters_dev` (device, LSG_COUNTER_WORDS u64) receives every integer result in one block so
 *  that a multi-GPU caller can sum-allreduce it; layout = lsg_counters. */
typedef struct {
  lsg_debug debug;
  lsg_confusion_out confusion;
  uint64_t events, spans, incidents, candidates, kept;
  uint64_t bad_idents; /* non-zero: some event's or span's ident was outside the identity table (or the caller's key_hint
                          did not hold) and the batch's results are void.  lsg_pipeline and lsg_pipeline_wait turn it into LSG_ERR_RANGE; a caller of
                          lsg_pipeline_dev checks it after reading the counters back (summed over ranks it stays non-zero) */
  uint64_t unsupported_events; /* events of an unsupported signal type (debug.unsupported_type = this x spans): what a caller
                          that shards a batch needs to rebuild the whole batch's Unmatched / UnsupportedType, which count
                          every (span, signal) pair, also those across shards (correlator.go:73-82) */
} lsg_counters;
#define LSG_COUNTER_WORDS (sizeof(lsg_counters) / 8)

typedef struct {
  const lsg_span *spans; size_t n_spans;
  const lsg_event *events; size_t n_events;
  const lsg_ident *idents; size_t n_idents;
  size_t n_incidents;
  const uint8_t *inc_actual;     /* [n_incidents] */
  const uint32_t *inc_expected;  /* [n_incidents] */
  const uint8_t *inc_expected_n; /* [n_incidents] */
  int64_t window_ns; double threshold; int fanout; double coverage_thr;
  /* outputs (device) */
  uint32_t *cand_event; uint8_t *cand_tier; double *attr_out; double *conf_out;
  double *inc_signals; double *posterior; uint8_t *top1;
  uint64_t *counters_dev;
  /* compact form of the per-span attributes (optional, NULL = not wanted): the kept candidates' signal index
   * (0..11, 0xff = empty slot) and value, [n_spans * fanout] each, in cand_event's order.  Without base attributes a
   * span's enriched attributes are exactly max(value) per signal over these (correlator.go:106-112) and its
   * confidence is the first candidate's tier confidence (correlator.go:113-118), so a caller that takes these may
   * pass attr_out = conf_out = NULL and move 42 instead of 119 bytes per span back to the host. */
  double *cand_value; uint8_t *cand_sig;
  /* optional statement about the batch (key_hint != 0): every span that can match (non-zero time, group != 0) has its
   * time inside [span_ts_lo, span_ts_hi], and no identity row's group exceeds max_grp_hint.  A replay knows both.  The
   * key layout then follows from the statement, nothing is read back and lsg_pipeline_dev never synchronises (it can
   * be captured in a CUDA graph once the workspace is warm).  The statement is checked on the device; if it does not
   * hold the batch's results are void and counters.bad_idents is non-zero. */
  uint32_t key_hint, max_grp_hint;
  int64_t span_ts_lo, span_ts_hi;
  /* optional: the events in their 24-byte upload form instead of `events` (which may then be NULL); n_events counts them.
   * Host pointer for lsg_pipeline / lsg_pipeline_submit, device pointer for lsg_pipeline_dev. */
  const lsg_event24 *events24; int64_t events24_base;
  /* optional (device for lsg_pipeline_dev, host otherwise): count / sum / log2 sketch per (incident, signal) of the
   * spans' enriched attributes, [n_incidents * 12], accumulated in the join's epilogue next to the maxima.  Only the
   * maxima (inc_signals) have a reference counterpart (correlator.go:109-111); these are this library's own features. */
  lsg_feature *features_out;
} lsg_pipeline_args;
/* asynchronous: returns after enqueueing on the context's stream (without key_hint it synchronises once inside, after
 * the first reduction, to read the key layout back) */
int lsg_pipeline_dev(lsg_ctx *ctx, const lsg_pipeline_args *args);
/* the same with HOST buffers in and out (every pointer of `args` is a host pointer; outputs may be NULL):
 * copies up, runs, copies back, synchronises.  What ProcessBatch + BuildAttributionsBayes +
 * BuildConfusionMatrix cost a Go caller end to end (processor.go:29, pipeline.go:31,65). */
int lsg_pipeline(lsg_ctx *ctx, const lsg_pipeline_args *args, lsg_counters *counters_out);

/* the same, double buffered: two slots (0, 1).  submit enqueues upload -> kernels -> download of one batch on three
 * streams and returns; wait blocks until that batch's results are in the caller's buffers.  With two batches in
 * flight the upload of one overlaps the kernels and the download of the other.  A slot's buffers must not be
 * touched between submit and wait and should be pinned. */
int lsg_pipeline_submit(lsg_ctx *ctx, const lsg_pipeline_args *args, int slot);
int lsg_pipeline_wait(lsg_ctx *ctx, int slot, lsg_counters *counters_out);

/* ---- a check of the JOIN-GROUP CONTRACT (top of this file) on a caller's batch: the library only pairs records of one
 *      group, so groups that are too fine lose matches silently.  Match() (dns.go:50-76) can accept a pair on the trace
 *      id, on the pod, or on the (service, node) pair; the contract holds exactly when each of those appears under one
 *      group.  out[0] = identity rows whose non-empty pod was first seen under another group, out[1] = rows whose
 *      non-empty (service, node) pair was, out[2] = spans / supported, timed events whose non-zero trace id was,
 *      out[3] = records whose ident is outside the identity table.  All zero = the contract holds; a non-zero count says
 *      it does not (how many records are counted depends on the order they are met in).  No reference counterpart. */
int lsg_validate_groups(lsg_ctx *ctx, const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                        const lsg_ident *idents, size_t n_idents, uint64_t out[4]);
int lsg_validate_groups_dev(lsg_ctx *ctx, const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                            const lsg_ident *idents, size_t n_idents, uint64_t out[4] /* HOST */);

/* ---- several GPUs of one box (no reference counterpart: the reference is one goroutine).  One context per
 *      device inside the calling process; the caller shards the batch so that no join group is split (shard by node /
 *      pod identity: Match() never pairs records of different groups, so every shard is a complete batch of its own) and
 *      passes one lsg_pipeline_args per rank with HOST buffers, exactly as for lsg_pipeline.  All ranks upload, compute
 *      and download at the same time; the one exchange is ncclAllReduce(ncclUint64, ncclSum) of the counter block.
 *      counters_out[r] (optional) = rank r's own counters, reduced_out = their sum (ratios such as accuracy are formed
 *      from the summed integers, so they do not depend on the reduction order) with debug.unmatched and
 *      debug.unsupported_type put on the whole batch's footing: the reference counts every (span, signal) pair of the
 *      batch (processor.go:29-49), also the pairs of a span and a signal that landed on different shards, which match
 *      nothing; so reduced_out equals what one context returns for the concatenated batch.  NCCL is looked up at run time
 *      (libnccl.so.2, or the path in LSG_NCCL_LIB) by lsg_multi_create; without it the call fails and the single-GPU
 *      entry points are unaffected.  Devices must be distinct.  Not thread-safe.  On failure of lsg_multi_create `*out`
 *      is still set when a handle exists: read lsg_multi_last_error, then lsg_multi_destroy. */
typedef struct lsg_multi lsg_multi;
int lsg_multi_create(const int *devices, int n_devices, lsg_multi **out);
void lsg_multi_destroy(lsg_multi *m);
const char *lsg_multi_last_error(const lsg_multi *m);
int lsg_multi_size(const lsg_multi *m);
lsg_ctx *lsg_multi_ctx(lsg_multi *m, int rank); /* rank r's context, e.g. for lsg_set_model; owned by the handle */
int lsg_multi_nccl_version(void);               /* NCCL_VERSION_CODE of the library found, 0 = none */
int lsg_multi_pipeline(lsg_multi *m, const lsg_pipeline_args *args /* n_devices */, lsg_counters *counters_out /* n_devices or NULL */,
                       lsg_counters *reduced_out);

/* ---- the 24-byte upload form (lsg_event24 above): host-side packing / unpacking of 32-byte records and the device pass
 *      that widens them.  lsg_pack_events24 returns the number of records whose ident / pid / conn / trace id does not
 *      fit the narrow fields (0 = the batch may go up in this form), < 0 on a bad argument.  No reference counterpart. */
long long lsg_pack_events24(const lsg_event *events, size_t n, int64_t base, lsg_event24 *out);
int lsg_unpack_events24(const lsg_event24 *in, size_t n, int64_t base, lsg_event *out);
int lsg_widen_events24_dev(lsg_ctx *ctx, const lsg_event24 *in, size_t n, int64_t base, lsg_event *out);

/* ---- per-kernel timing with CUDA events on the context's stream (bench.py's roofline numbers) ---------
 *  lsg_profile_read synchronises the stream, adds the finished intervals to per-kernel totals and copies
 *  up to `cap` of them out (milliseconds, launches-groups counted); returns the number of kinds. */
/* key layout of the most recent event sort: total key bits, log2 of the time quantum in ns, radix passes */
int lsg_last_sort_info(const lsg_ctx *ctx, uint32_t *key_bits, uint32_t *time_shift, uint32_t *passes);
int lsg_profile_enable(lsg_ctx *ctx, int on);
const char *lsg_profile_name(int kind);
int lsg_profile_read(lsg_ctx *ctx, double *ms, uint64_t *count, int cap, int reset);

/* ---- rows either side of the path ---------------------------------------------------------
 *
 * F2  lsg_decode_ringbuf: the 40-byte packed little-endian struct llm_slo_event (ebpf/c/llm_slo_event.h:32-42,
 *     bpfEvent in pkg/collector/ringbuf.go:42-52) -> lsg_event, doing what decodeBPFEvent + toProbeEvent +
 *     signalFromType + convertValue do per record (ringbuf.go:152-238): kernel signal type 1..9 -> signal index,
 *     ns -> ms except for counts and cpu_steal, ConnTuple present when a port is non-zero.  The metadata the agent
 *     stamps on every event of one ring buffer (pod/node/service -> `ident`, trace id) are call arguments;
 *     ts = timestamp_ns + ts_offset_ns (the reference stamps time.Now() at decode time, ringbuf.go:172, which a
 *     replay cannot reproduce; pass the boot-time offset).  conn is a 32-bit mix of (src port, dst port, dst ip).
 *     status_out (optional) is F4's code; errno_out (optional) the raw errno.
 * F4  lsg_signal_status: signalStatus (pkg/signals/generator.go:203-242): 0 ok, 1 warning, 2 error.
 * F1  lsg_retry_storm: RetryStormDetector.Record (pkg/correlation/retry_storm.go:46-62) replayed over a batch:
 *     for every tcp_retransmits_total event, the number of retransmit events of the same pod with a time in
 *     [ts - window, ts] that arrived no later, and whether that reaches `threshold`; 0 for other events.  Equals
 *     the reference when each pod's events arrive in time order (its prefix-only pruning, retry_storm.go:98-110,
 *     makes out-of-order arrivals order-dependent). */
int lsg_decode_ringbuf(lsg_ctx *ctx, const void *raw /* n*40 bytes */, size_t n, uint32_t ident, uint32_t trace,
                       int64_t ts_offset_ns, lsg_event *events_out, uint8_t *status_out, int32_t *errno_out);
int lsg_decode_ringbuf_dev(lsg_ctx *ctx, const void *raw, size_t n, uint32_t ident, uint32_t trace, int64_t ts_offset_ns,
                           lsg_event *events_out, uint8_t *status_out, int32_t *errno_out);
int lsg_signal_status(lsg_ctx *ctx, const lsg_event *events, size_t n, uint8_t *status_out);
int lsg_signal_status_dev(lsg_ctx *ctx, const lsg_event *events, size_t n, uint8_t *status_out);
int lsg_retry_storm(lsg_ctx *ctx, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                    int64_t window_ns, int threshold, uint32_t *count_out, uint8_t *storm_out);
int lsg_retry_storm_dev(lsg_ctx *ctx, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                        int64_t window_ns, int threshold, uint32_t *count_out, uint8_t *storm_out);

/* F4  lsg_segment_quantiles: quantile() of pkg/slo/calculator.go:121-149 (same body: pkg/releasegate/gate.go:737-760)
 *     for many value lists in one call, e.g. one list per (scenario, phase, signal) for the `*_p95` / `*_p50`
 *     assertions of test/incident-lab/scenarios/<name>.yaml, or per node for releasegate's node p95.  Record i belongs
 *     to list segment[i]; ids >= n_segments are ignored; segment == NULL means one list (n_segments must be <= 1).
 *     out[s * n_q + j] = quantile(list s, q[j]): 0 for an empty list, q clamped to [0, 1], a sorted copy with NaN
 *     first (sort.Float64s), pos = q * (len - 1), linear interpolation between floor(pos) and ceil(pos) in the
 *     reference's operation order.  count_out[s] (optional) = len(list s).  n < 2^32.  NaN in q is rejected by
 *     the host variant.  The same kernels give the north star's "quantile sketch" feature an exact meaning. */
int lsg_segment_quantiles(lsg_ctx *ctx, const double *values, const uint32_t *segment, size_t n, size_t n_segments,
                          const double *q, size_t n_q, double *out /* n_segments * n_q */, uint64_t *count_out);
int lsg_segment_quantiles_dev(lsg_ctx *ctx, const double *values, const uint32_t *segment, size_t n, size_t n_segments,
                              const double *q, size_t n_q, double *out, uint64_t *count_out);

/* ---- synthetic workload: record i is a pure function of (seed, i), generated on
 *      the device by lsg_synth_dev and on the host by lsg_synth_host with identical results. */
typedef struct {
  uint64_t seed;
  uint64_t n_events, n_spans;
  uint32_t n_pods, pods_per_node, pods_per_group;
  uint32_t spans_per_incident;
  uint32_t multi_fault;        /* 0: single-fault incidents; 1: 2-3 concurrent faults               */
  int64_t t0_ns, horizon_ns;
  uint32_t rank, world;        /* shard: this rank's pods are [rank*n_pods, (rank+1)*n_pods) of the global space */
} lsg_synth_cfg;
size_t lsg_synth_n_idents(const lsg_synth_cfg *cfg);
size_t lsg_synth_n_incidents(const lsg_synth_cfg *cfg);
int lsg_synth_dev(lsg_ctx *ctx, const lsg_synth_cfg *cfg, lsg_span *spans, lsg_event *events, lsg_ident *idents,
                  uint8_t *inc_actual, uint32_t *inc_expected, uint8_t *inc_expected_n);
int lsg_synth_host(const lsg_synth_cfg *cfg, lsg_span *spans, lsg_event *events, lsg_ident *idents,
                   uint8_t *inc_actual, uint32_t *inc_expected, uint8_t *inc_expected_n);

#ifdef __cplusplus
}
#endif
#endif /* LLMSLO_GPU_H */
