// common.cuh — context, workspace, error plumbing and device helpers shared by the kernels of
// libllmslo_gpu.so.  sm_100a only; no CPU fallback anywhere in this library.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/llmslo_gpu.h"

// ---- workspace slots (grow-only device buffers owned by the context) ---------------------------
enum {
  WS_EVENTS_IN = 0,  // host-API staging: events
  WS_SPANS_IN,       // host-API staging: spans
  WS_IDENTS,         // host-API staging: identity table
  WS_KEYS_A,         // radix sort: key ping
  WS_KEYS_B,         // radix sort: key pong
  WS_IDX_A,          // radix sort: index ping
  WS_IDX_B,          // radix sort: index pong
  WS_HIST,           // radix sort: per-tile digit histograms
  WS_SMALL,          // reductions / counters / digit bases (a few KB)
  WS_EV_SORTED,      // events in (grp, time) order
  WS_SPAN_KEYS_A,
  WS_SPAN_KEYS_B,
  WS_SPAN_IDX_A,
  WS_SPAN_IDX_B,
  WS_OUT0,           // host-API staging of outputs
  WS_OUT1,
  WS_OUT2,
  WS_OUT3,
  WS_OUT4,
  WS_OUT5,
  WS_OUT6,
  WS_OUT7,
  WS_IN0,            // host-API staging of misc inputs
  WS_IN1,
  WS_IN2,
  WS_IN3,
  WS_IN4,
  WS_FEAT,           // incident feature accumulators
  WS_RANGES,         // per-span event ranges of the join
  WS_SPAN_HIST,      // radix sort of the spans: per-tile digit histograms
  WS_BINS,           // counting sort: one counter per key
  WS_SCAN,           // counting sort: block sums of the scan
  WS_IDENT_PC,       // join: (pod, service+node class) per identity row
  WS_IDENT_TABLE,    // join: hash table that numbers the (service, node) classes
  WS_BK_EV,          // bucket partition: events grouped by (group, time slab) bucket
  WS_BK_EV_ORIG,     // ... their input indices
  WS_BK_EV_ENDS,     // ... bucket ends (and, before the scatter, counts / starts)
  WS_BK_SP,          // the same three for the spans
  WS_BK_SP_ORIG,
  WS_BK_SP_ENDS,
  WS_BK_TMP,         // two-level partition: records grouped by super-bucket
  WS_BK_TMP_ORIG,
  WS_BK_SUPER,       // two-level partition: super-bucket cursors
  WS_TILES,          // slab join: tile list
  WS_EV_WIDE,        // events widened from the 24-byte upload form
  WS_MODEL_TABLE,    // attribution: posteriors and label of every elevated-signal mask under the current model
  WS_IN5,            // host-API staging: feature output
  WS_PIPE,           // first of 2 x PIPE_BUFS staging buffers of the double-buffered host pipeline
  WS_PIPE_END = WS_PIPE + 2 * 20 - 1,
  WS_COUNT
};

// device copy of the model: logs are taken on the host (std::log, the oracle's libm) so that the
// per-domain sums are bit-identical to the CPU restatement; the kernel only adds (bayesian.go:230-242).
struct DevModel {
  double log_prior[LSG_NDOM];          // log(prior or 1e-10) + log_extra
  double log_elev[LSG_NSIG][LSG_NDOM];  // log clamp(p)      (0 if the signal is not in Likelihoods)
  double log_not[LSG_NSIG][LSG_NDOM];   // log clamp(1 - p)
  double log_extra[LSG_NDOM];
  double thresh[LSG_NSIG];
  uint32_t sig_mask;
};

#ifdef __CUDACC__
// doubles as order-preserving unsigned keys, so that max is an integer atomicMax; key 0 is never produced by a
// non-NaN value and stands for "absent"
__device__ __forceinline__ unsigned long long f64_key(double v) {
  unsigned long long b = (unsigned long long)__double_as_longlong(v);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double key_f64(unsigned long long k) {
  if (k == 0) return __longlong_as_double(0x7ff8000000000000ll);
  unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  return __longlong_as_double((long long)b);
}
#endif

// ---- sort key: (join group, quantised time) ------------------------------------------------------
// live key = (grp << tbits) | ((ts - ts_min) >> shift); records that can match nothing (group 0, zero
// Time, unsupported signal type) get the key (max_grp + 1) << tbits and sort behind everything a span
// ever searches.
struct KeySpec {
  int64_t ts_min;   // origin of the time axis; events before it can match nothing
  int64_t ts_hi;    // events after it can match nothing (last span + window, or the last live event)
  uint32_t shift;   // time quantum = 2^shift ns
  uint32_t tbits;   // bits of the quantised time
  uint32_t bits;    // total key bits (gbits + tbits)
  uint64_t dead;    // key of records that can match nothing: one group past the last real one, so they sort last
};

// per-kernel timing (bench.py's roofline): CUDA events recorded on the context's stream around each launch
enum {
  PK_STATS = 0, PK_KEYS, PK_HIST, PK_SCAN, PK_SCATTER, PK_GATHER, PK_SPAN_SORT, PK_JOIN, PK_FEATURES, PK_ATTRIBUTE,
  PK_CONFUSION, PK_PAIRS, PK_RANGES, PK_BK_COUNT, PK_BK_SCATTER, PK_BK_SPANS, PK_TILES, PK_WIDEN, PK_COUNT
};
struct ProfSpan { int kind; cudaEvent_t a, b; };

struct lsg_ctx {
  bool prof = false;
  std::vector<ProfSpan> prof_open;      // recorded, not yet collected
  std::vector<cudaEvent_t> prof_pool;   // reusable events
  double prof_ms[PK_COUNT] = {};
  uint64_t prof_n[PK_COUNT] = {};
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  cudaStream_t side = nullptr;          // the span sort runs here, next to the event sort
  cudaStream_t up = nullptr, down = nullptr;                    // lsg_pipeline_submit: uploads / downloads next to the kernels
  cudaEvent_t ev_up[2] = {}, ev_done[2] = {}, ev_down[2] = {};  // per slot: inputs on the device, kernels finished, results on the host
  bool slot_busy[2] = {false, false};
  void *slot_counters_host[2] = {};     // pinned, sizeof(lsg_counters) each
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  int sm_count = 148;
  std::string err;
  uint64_t launches = 0;
  void *ws[WS_COUNT] = {};
  size_t ws_cap[WS_COUNT] = {};
  void *pinned = nullptr;  // small pinned readback buffer
  size_t pinned_cap = 0;
  DevModel model;
  bool model_set = false;
  bool model_table_valid = false;   // WS_MODEL_TABLE matches `model`
  KeySpec last_key = {};   // key layout of the most recent event sort (diagnostics)
  int join_path = -1;      // -1: LSG_JOIN_PATH decides (default perm), 0: permutation sort + gathering join, 1: bucket partition + slab join,
                           // 2 / 3: as 0 with the sliding-tile / the fixed-chunk form of the tile kernel forced
};

int lsg_fail(lsg_ctx *ctx, int code, const char *fmt, ...);
int lsg_ws(lsg_ctx *ctx, int slot, size_t bytes, void **out);

#define LSG_CUDA(ctx, call)                                                                            \
  do {                                                                                                 \
    cudaError_t e__ = (call);                                                                          \
    if (e__ != cudaSuccess)                                                                            \
      return lsg_fail(ctx, LSG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)
#define LSG_TRY(expr)          \
  do {                         \
    int rc__ = (expr);         \
    if (rc__ != 0) return rc__; \
  } while (0)
#define LSG_LAUNCH_CHECK(ctx)      \
  do {                             \
    (ctx)->launches++;             \
    LSG_CUDA(ctx, cudaGetLastError()); \
  } while (0)

// scoped timer: records an event pair around the launches issued while it is alive (only when profiling)
struct ProfScope {
  lsg_ctx *ctx; int kind; cudaEvent_t a = nullptr, b = nullptr;
  ProfScope(lsg_ctx *c, int k) : ctx(c), kind(k) {
    if (!ctx->prof) return;
    auto take = [&]() { cudaEvent_t e = nullptr; if (!ctx->prof_pool.empty()) { e = ctx->prof_pool.back(); ctx->prof_pool.pop_back(); } else cudaEventCreate(&e); return e; };
    a = take(); b = take();
    cudaEventRecord(a, ctx->stream);
  }
  ~ProfScope() {
    if (!a) return;
    cudaEventRecord(b, ctx->stream);
    ctx->prof_open.push_back(ProfSpan{kind, a, b});
  }
};

// ---- device helpers ---------------------------------------------------------------------------
#define LSG_HD __host__ __device__ __forceinline__

static constexpr int64_t kMs = 1000000;

// |a - b| as an exact unsigned magnitude (dns.go:111 compares |float64(a.Sub(b))| <= float64(window);
// identical for window < 2^53 ns, which the API enforces).
LSG_HD uint64_t lsg_absdiff(int64_t a, int64_t b) { return a > b ? (uint64_t)a - (uint64_t)b : (uint64_t)b - (uint64_t)a; }

// correlation.Match (dns.go:50-76) on interned records, given |dt| and that neither time is zero and
// |dt| <= window already holds.  si/ei are the identity rows.  Returns the tier.
LSG_HD int lsg_match_tier(uint64_t adt, uint32_t s_trace, uint32_t s_pid, uint32_t s_conn, const lsg_ident &si,
                          uint32_t e_trace, uint32_t e_pid, uint32_t e_conn, const lsg_ident &ei) {
  if (s_trace != 0 && s_trace == e_trace) return LSG_TIER_TRACE;                       // dns.go:58-60
  const bool same_pod = si.pod != 0 && si.pod == ei.pod;
  if (same_pod && s_pid > 0 && s_pid == e_pid && adt <= (uint64_t)(100 * kMs)) return LSG_TIER_POD_PID;    // dns.go:61-64
  if (same_pod && s_conn != 0 && s_conn == e_conn && adt <= (uint64_t)(250 * kMs)) return LSG_TIER_POD_CONN;  // dns.go:65-68
  if (si.service != 0 && si.service == ei.service && si.node != 0 && si.node == ei.node &&
      adt <= (uint64_t)(500 * kMs))
    return LSG_TIER_SVC_NODE;                                                          // dns.go:69-73
  return LSG_TIER_NONE;
}

LSG_HD double lsg_tier_conf(int tier) {
  return tier == 1 ? 1.0 : tier == 2 ? 0.9 : tier == 3 ? 0.8 : tier == 4 ? 0.65 : 0.0;
}

#ifdef __CUDACC__
// 128-bit streaming loads/stores of the 32-byte records (two per record)
__device__ __forceinline__ uint4 ld_nc_v4(const void *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_v4(const void *p) { return *reinterpret_cast<const uint4 *>(p); }
__device__ __forceinline__ void st_v4(void *p, uint4 v) { *reinterpret_cast<uint4 *>(p) = v; }

// 256-bit loads/stores (sm_100: LDG.E.256 / STG.E.256): one request per 32-byte record.  Needs 32-byte alignment,
// which every record array has (cudaMalloc / torch allocations; documented in the header for caller buffers).
struct U8 { uint4 a, b; };
__device__ __forceinline__ U8 ld_nc_v8(const void *p) {
  unsigned long long x, y, z, w;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(x), "=l"(y), "=l"(z), "=l"(w) : "l"(p));
  U8 r;
  r.a = make_uint4((uint32_t)x, (uint32_t)(x >> 32), (uint32_t)y, (uint32_t)(y >> 32));
  r.b = make_uint4((uint32_t)z, (uint32_t)(z >> 32), (uint32_t)w, (uint32_t)(w >> 32));
  return r;
}
__device__ __forceinline__ void st_v8(void *p, const U8 v) {
  const unsigned long long x = ((unsigned long long)v.a.y << 32) | v.a.x, y = ((unsigned long long)v.a.w << 32) | v.a.z;
  const unsigned long long z = ((unsigned long long)v.b.y << 32) | v.b.x, w = ((unsigned long long)v.b.w << 32) | v.b.z;
  asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(x), "l"(y), "l"(z), "l"(w) : "memory");
}

struct EventRec {  // register image of an lsg_event
  int64_t ts;
  double value;
  uint32_t ident, pid, conn, trace_sig;
};
__device__ __forceinline__ EventRec load_event(const lsg_event *p) {
  const U8 u = ld_nc_v8(p);
  const uint4 a = u.a, b = u.b;
  EventRec r;
  r.ts = (int64_t)(((uint64_t)a.y << 32) | a.x);
  r.value = __longlong_as_double((long long)(((uint64_t)a.w << 32) | a.z));
  r.ident = b.x; r.pid = b.y; r.conn = b.z; r.trace_sig = b.w;
  return r;
}
struct SpanRec {
  int64_t ts;
  uint32_t ident, pid, conn, trace, incident;
};
__device__ __forceinline__ SpanRec load_span(const lsg_span *p) {
  const U8 u = ld_nc_v8(p);
  const uint4 a = u.a, b = u.b;
  SpanRec r;
  r.ts = (int64_t)(((uint64_t)a.y << 32) | a.x);
  r.ident = a.z; r.pid = a.w; r.conn = b.x; r.trace = b.y; r.incident = b.z;
  return r;
}
__device__ __forceinline__ lsg_ident load_ident(const lsg_ident *t, uint32_t row) {
  uint4 a = __ldg(reinterpret_cast<const uint4 *>(t + row));
  lsg_ident r;
  r.pod = a.x; r.node = a.y; r.service = a.z; r.grp = a.w;
  return r;
}
#endif

LSG_HD uint64_t lsg_make_key(const KeySpec &k, uint32_t grp, int64_t ts) {
  return ((uint64_t)grp << k.tbits) | (((uint64_t)ts - (uint64_t)k.ts_min) >> k.shift);
}


// stage kernels (defined across the .cu files)
int lsg_launch_match_pairs(lsg_ctx *, const lsg_span *, const lsg_event *, size_t, const lsg_ident *, size_t, int64_t, double,
                           const uint8_t *, const uint8_t *, uint8_t *, uint8_t *, uint64_t *eval_dev);
struct EventColumns {     // events grouped by key as columns (what the join stages into shared memory)
  const int64_t *ts;
  const double *val;
  const uint32_t *ident, *pid, *conn, *trsig;
};
// what one look at every event (and span) of a batch collects, in device memory at the start of WS_SMALL
struct LsgBatchStats {
  long long ts_min, ts_max;                 // over the live events, or over the live spans (see lsg_sort_events_prepare)
  unsigned long long n_live, n_unsupported; // events; n_live only when the events were looked at before the sort
  unsigned int max_grp, bad;                // bad: bit 0 an event's, bit 1 a span's ident is outside the identity table,
                                            // bit 2 the caller's key hint does not hold
  unsigned int pad[2];
};

// what a caller may state about a batch so that nothing has to be read back before the sort (lsg_pipeline_args.key_hint)
struct LsgKeyHint {
  long long span_ts_lo, span_ts_hi;   // every live span's time lies in [lo, hi]
  unsigned int max_grp;               // no identity row has a larger group
};

struct SortedEvents {
  EventColumns col;       // set when the records were not asked for (the join's input)
  const lsg_event *rec;   // set when the caller asked for records (lsg_sort_events_dev)
  const uint32_t *perm;   // position -> input index
  const void *keys;       // radix path: the sorted keys (uint32_t if key.bits <= 31 else uint64_t); NULL on the counting path
  const uint32_t *bins;   // counting path: bins[k] = start of key k's run, bins[k + 1] its end; NULL on the radix path
  KeySpec key;
  size_t n;               // all events
  size_t n_live;          // upper bound of the events that can match (they come first); exact on the radix path only
                          // when the key range came from the events
  const LsgBatchStats *stats_dev;   // device copy, complete once the sort kernels ran (n_unsupported, bad)
  bool tally_in_sort;               // the sort's first pass adds n_unsupported / bad to stats_dev (key range from the spans)
};
// sorted_out NULL: records into the context workspace.  `exact`: stable, nanosecond keys, always the radix path.
int lsg_sort_events(lsg_ctx *, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents, int64_t window_ns,
                    bool exact, lsg_event *sorted_out, SortedEvents *out);
// With `spans` the key range comes from the spans (first live span - window .. last live span + window: an event
// outside it can match nothing and is treated like a dead record), so the events are not read before the sort; the
// unsupported-type count and the events' ident check then come out of the sort's own first pass (stats_dev).
// Without spans (the sort as its own entry point) one pass over the events sets the range.
int lsg_sort_events_prepare(lsg_ctx *, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                            int64_t window_ns, bool exact, SortedEvents *out, const lsg_span *spans, size_t n_spans,
                            const LsgKeyHint *hint);
// want_copy false (and sorted_out NULL): only the permutation (and bins / keys) is produced, no sorted copy of the events
int lsg_sort_events_finish(lsg_ctx *, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                           lsg_event *sorted_out, SortedEvents *out, bool allow_counting, bool want_copy);
int lsg_sort_signal_by_pod(lsg_ctx *, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents, uint32_t sig,
                           const uint64_t **keys_out, const uint32_t **perm_out, KeySpec *ks_out, size_t *n_live_out);
int lsg_sort_spans(lsg_ctx *, const lsg_span *spans, size_t n, const lsg_ident *idents, size_t n_idents, const KeySpec &key,
                   const uint32_t **perm_out);
// ---- bucket partition (sort.cu) ---------------------------------------------------------------------
// The 32-byte records are moved ONCE into coarse buckets: bucket = sort key >> LSG_BUCKET_SHIFT, i.e. one join group
// x 64 consecutive time quanta (longer than the window, at most twice as long).  Inside a bucket the order is
// whatever the cursor atomics gave; the slab join orders a tile by time quantum in shared memory when it stages it.
constexpr uint32_t LSG_BUCKET_SHIFT = 6;
struct Bucketed {
  const void *rec;         // lsg_event / lsg_span records grouped by bucket
  const uint32_t *orig;    // position -> input index
  const uint32_t *ends;    // ends[b] = one past the last record of bucket b; ends[-1] = 0 is readable
  uint32_t n_buckets;      // buckets of records that can match; for spans bucket n_buckets holds the ones that cannot
  KeySpec key;
};
// key layout of the bucket path derived from the sort's (coarser quantum for sparse batches, tbits >= bucket shift)
KeySpec lsg_bucket_keyspec(const KeySpec &ks, size_t n_events, int64_t window_ns);
bool lsg_bucket_path_ok(const KeySpec &bk, size_t n_events, size_t n_spans, int64_t window_ns, int fanout);
// stats_dev: the first pass adds the unsupported-type count and the ident check of the events (NULL: not wanted)
int lsg_bucket_events(lsg_ctx *, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents, const KeySpec &bk,
                      LsgBatchStats *stats_dev, Bucketed *out);
int lsg_bucket_spans(lsg_ctx *, const lsg_span *spans, size_t n, const lsg_ident *idents, size_t n_idents, const KeySpec &bk,
                     Bucketed *out);

struct JoinArgs {
  const lsg_span *spans; size_t n_spans; const uint32_t *span_perm;
  SortedEvents ev; const lsg_event *events_in; const lsg_ident *idents; size_t n_idents;
  int64_t window_ns; double threshold; int fanout;
  const double *attr_base; uint32_t *cand_event; uint8_t *cand_tier; double *attr_out; double *conf_out;
  double *retrieval_out; uint32_t *n_cand_out;
  uint64_t *counters;  // device: [0] matched pairs, [1] low_confidence, [2] fanout_dropped, [3] candidates, [4] kept
  // optional fusion of the A7 glue: per (incident, signal) max keys updated straight from the join's epilogue
  unsigned long long *inc_keys; size_t n_incidents;
  const uint2 *ident_pc;   // lsg_ident_classes() of `idents`, or NULL: lsg_launch_join computes it
  double *cand_value; uint8_t *cand_sig;   // optional: value and signal of every kept candidate, [n_spans * fanout]
  lsg_feature *feat = nullptr;             // optional: count / sum / sketch per (incident, signal), zeroed by the caller
};
int lsg_launch_join(lsg_ctx *, const JoinArgs &);
// the slab join: events and spans bucketed by lsg_bucket_*; the JoinArgs fields about the sorted events (ev, span_perm)
// are not used, events_in / spans are the caller's arrays (outputs are indexed by input position)
int lsg_launch_join_slab(lsg_ctx *, const JoinArgs &, const Bucketed &ev, const Bucketed &sp);
// per identity row: x = pod id, y = class of its (service, node) pair (0 when either is empty; equal pairs get
// equal classes), so that the join compares two words per (span, event) pair instead of loading identity rows
int lsg_ident_classes(lsg_ctx *, const lsg_ident *idents, size_t n_idents, const uint2 **out);
int lsg_launch_widen_events24(lsg_ctx *, const lsg_event24 *in, size_t n, int64_t base, lsg_event *out);
int lsg_launch_incident_finish(lsg_ctx *, const unsigned long long *keys, size_t n_incidents, double *signals_out);
int lsg_launch_incident_features(lsg_ctx *, const double *span_attr, const uint32_t *span_incident, const lsg_span *spans,
                                 size_t n_spans, size_t n_incidents, double *signals_out, lsg_feature *features_out);
int lsg_launch_attribute(lsg_ctx *, const double *signals, const uint8_t *has_signals, const uint8_t *rule_label, size_t n,
                         double *posterior, uint8_t *top1, uint16_t *elev_mask);
int lsg_launch_confusion(lsg_ctx *, const uint8_t *actual, const uint8_t *predicted, const uint32_t *expected,
                         const uint8_t *expected_n, const double *posterior, double coverage_thr, double hyp_floor, size_t n,
                         uint64_t *out_dev /* lsg_confusion_out layout, zeroed by the caller */);
