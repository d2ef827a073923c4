// api.cu — the C ABI of include/llmslo_gpu.h: context, workspace, host-buffer wrappers and the fused
// device-resident pipeline.  Everything that computes is a CUDA kernel in the sibling .cu files; this
// file only moves buffers, validates arguments and launches.  No CPU fallback.
#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <new>

#include "common.cuh"

static_assert(sizeof(lsg_event) == 32 && sizeof(lsg_span) == 32 && sizeof(lsg_ident) == 16, "record layout");
static_assert(sizeof(lsg_feature) == 80, "feature layout");
static_assert(sizeof(lsg_event24) == 24, "upload record layout");

int lsg_fail(lsg_ctx *ctx, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf;
  return code;
}

int lsg_ws(lsg_ctx *ctx, int slot, size_t bytes, void **out) {
  if (bytes == 0) bytes = 16;
  if (ctx->ws_cap[slot] < bytes) {
    if (ctx->ws[slot]) {
      cudaStreamSynchronize(ctx->stream);
      cudaFree(ctx->ws[slot]);
      ctx->ws[slot] = nullptr;
      ctx->ws_cap[slot] = 0;
    }
    size_t cap = bytes + bytes / 8;   // headroom so that slowly growing batches do not reallocate every call
    cap = (cap + 255) & ~(size_t)255;
    cudaError_t e = cudaMalloc(&ctx->ws[slot], cap);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return lsg_fail(ctx, LSG_ERR_NOMEM, "cudaMalloc of %zu bytes (workspace slot %d) failed: %s", cap, slot, cudaGetErrorString(e));
    }
    ctx->ws_cap[slot] = cap;
  }
  *out = ctx->ws[slot];
  return 0;
}

namespace {

constexpr size_t SMALL_JOIN = 256, SMALL_EVAL = 512, SMALL_CONF = 1024;

void build_dev_model(const lsg_model *m, DevModel *d) {
  auto clampl = [](double p) { return p < 0.01 ? 0.01 : (p > 0.99 ? 0.99 : p); };   // bayesian.go:303-311
  for (int k = 0; k < LSG_NDOM; ++k) {
    double prior = m->prior[k];
    if (!(prior > 0)) prior = 1e-10;   // bayesian.go:231-233
    d->log_prior[k] = std::log(prior);
    d->log_extra[k] = m->log_extra[k];
  }
  for (int s = 0; s < LSG_NSIG; ++s) {
    d->thresh[s] = m->thresh[s];
    for (int k = 0; k < LSG_NDOM; ++k) {
      const double p = m->lik[s][k];
      if (std::isnan(p)) {   // likelihoodFor: domain missing -> 0.5 (bayesian.go:293-296)
        d->log_elev[s][k] = d->log_not[s][k] = std::log(0.5);
      } else {
        d->log_elev[s][k] = std::log(clampl(p));
        d->log_not[s][k] = std::log(clampl(1 - p));
      }
    }
  }
  d->sig_mask = m->sig_mask;
}

int set_device(lsg_ctx *ctx) {
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  return 0;
}

template <typename T>
int h2d(lsg_ctx *ctx, int slot, const T *host, size_t n, const T **dev) {
  *dev = nullptr;
  if (!host) return 0;
  void *p;
  LSG_TRY(lsg_ws(ctx, slot, n * sizeof(T), &p));
  if (n) LSG_CUDA(ctx, cudaMemcpyAsync(p, host, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  *dev = (const T *)p;
  return 0;
}
template <typename T>
int dalloc(lsg_ctx *ctx, int slot, bool want, size_t n, T **dev) {
  *dev = nullptr;
  if (!want) return 0;
  void *p;
  LSG_TRY(lsg_ws(ctx, slot, n * sizeof(T), &p));
  *dev = (T *)p;
  return 0;
}
template <typename T>
int d2h(lsg_ctx *ctx, T *host, const T *dev, size_t n) {
  if (host && dev && n) LSG_CUDA(ctx, cudaMemcpyAsync(host, dev, n * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  return 0;
}

int norm_params(lsg_ctx *ctx, int64_t *window_ns, double *threshold, int *fanout) {
  if (*window_ns <= 0) *window_ns = 2000 * kMs;            // dns.go:51-53
  if (*threshold <= 0) *threshold = 0.7;                  // correlator.go:58-61
  if (fanout && *fanout <= 0) *fanout = 3;                // correlator.go:62-66
  if (*window_ns > LSG_MAX_WINDOW_NS) return lsg_fail(ctx, LSG_ERR_INVALID, "window_ns %lld exceeds 2^53-1", (long long)*window_ns);
  if (fanout && *fanout > LSG_MAX_FANOUT) return lsg_fail(ctx, LSG_ERR_INVALID, "fanout %d exceeds LSG_MAX_FANOUT=%d", *fanout, LSG_MAX_FANOUT);
  return 0;
}

// sort + join on device pointers; counters_host[0..4] = matched, low_confidence, fanout_dropped, candidates, kept
int process_batch_core(lsg_ctx *ctx, const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                       const lsg_ident *idents, size_t n_idents, int64_t window_ns, double threshold, int fanout,
                       const double *attr_base, uint32_t *cand_event, uint8_t *cand_tier, double *attr_out, double *conf_out,
                       double *retrieval_out, uint32_t *n_cand_out, uint64_t **join_counters_dev, SortedEvents *sorted,
                       unsigned long long *inc_keys = nullptr, size_t n_incidents = 0, double *cand_value = nullptr,
                       uint8_t *cand_sig = nullptr, const LsgKeyHint *hint = nullptr, lsg_feature *feat = nullptr,
                       const int64_t *e24_base = nullptr) {
  LSG_TRY(lsg_sort_events_prepare(ctx, events, n_events, idents, n_idents, window_ns, false, sorted, spans, n_spans, hint));
  if (e24_base) {
    // events that came up in the 24-byte form: their time offsets only reach from `base` to base + 2^42 - 2, which must
    // cover everything a span can match (the key's time axis: first span - window .. last span + window)
    const KeySpec &k = sorted->key;
    if (k.ts_min <= k.ts_hi && (*e24_base > k.ts_min || (uint64_t)k.ts_hi - (uint64_t)*e24_base >= (((uint64_t)1 << LSG_E24_TS_BITS) - 1)))
      return lsg_fail(ctx, LSG_ERR_RANGE, "events24_base %lld does not cover the spans' time range widened by the window "
                      "[%lld, %lld] with a %d-bit offset", (long long)*e24_base, (long long)k.ts_min, (long long)k.ts_hi,
                      LSG_E24_TS_BITS);
  }
  void *small_v;
  LSG_TRY(lsg_ws(ctx, WS_SMALL, 4096, &small_v));
  // Two ways from here.  Default: permutation sort + the join that gathers its own tiles (the faster one as measured,
  // profiles/r02_*).  The other: the records are moved once into (group, time slab) buckets and the slab join stages
  // contiguous tiles (sort.cu "bucket partition", join_slab.cuh): DRAM traffic close to the algorithmic bytes, but the
  // in-tile counting sort and its barriers cost more than the gather's latency.  lsg_set_join_path / LSG_JOIN_PATH=slab.
  static const char *path_env = getenv("LSG_JOIN_PATH");
  const bool want_slab = ctx->join_path == 1 || (ctx->join_path < 0 && path_env && strcmp(path_env, "slab") == 0);
  const KeySpec bk = lsg_bucket_keyspec(sorted->key, n_events, window_ns);
  if (want_slab && lsg_bucket_path_ok(bk, n_events, n_spans, window_ns, fanout)) {
    Bucketed bev, bsp;
    const uint2 *ident_pc = nullptr;
    {
      LSG_CUDA(ctx, cudaEventRecord(ctx->ev_fork, ctx->stream));
      LSG_CUDA(ctx, cudaStreamWaitEvent(ctx->side, ctx->ev_fork, 0));
      cudaStream_t main_stream = ctx->stream;
      ctx->stream = ctx->side;   // identity classes and the spans' partition run next to the events' partition
      int rc = lsg_ident_classes(ctx, idents, n_idents, &ident_pc);
      if (!rc) rc = lsg_bucket_spans(ctx, spans, n_spans, idents, n_idents, bk, &bsp);
      ctx->stream = main_stream;
      if (rc) return rc;
      LSG_CUDA(ctx, cudaEventRecord(ctx->ev_join, ctx->side));
    }
    LSG_TRY(lsg_bucket_events(ctx, events, n_events, idents, n_idents, bk, const_cast<LsgBatchStats *>(sorted->stats_dev), &bev));
    LSG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
    ctx->last_key = bk;
    JoinArgs ja;
    memset(&ja, 0, sizeof ja);
    ja.spans = spans; ja.n_spans = n_spans;
    ja.ev = *sorted; ja.events_in = events; ja.idents = idents; ja.n_idents = n_idents;
    ja.window_ns = window_ns; ja.threshold = threshold; ja.fanout = fanout;
    ja.attr_base = attr_base; ja.cand_event = cand_event; ja.cand_tier = cand_tier; ja.attr_out = attr_out;
    ja.conf_out = conf_out; ja.retrieval_out = retrieval_out; ja.n_cand_out = n_cand_out;
    ja.counters = (uint64_t *)((char *)small_v + SMALL_JOIN);
    ja.inc_keys = inc_keys; ja.n_incidents = n_incidents;
    ja.ident_pc = ident_pc;
    ja.cand_value = cand_value; ja.cand_sig = cand_sig;
    ja.feat = feat;
    LSG_TRY(lsg_launch_join_slab(ctx, ja, bev, bsp));
    *join_counters_dev = ja.counters;
    return 0;
  }
  // the span sort only needs the key layout: it runs on the side stream while the event sort runs on the main one
  const uint32_t *span_perm = nullptr;
  const uint2 *ident_pc = nullptr;
  {
    LSG_CUDA(ctx, cudaEventRecord(ctx->ev_fork, ctx->stream));
    LSG_CUDA(ctx, cudaStreamWaitEvent(ctx->side, ctx->ev_fork, 0));
    cudaStream_t main_stream = ctx->stream;
    ctx->stream = ctx->side;
    int rc = lsg_ident_classes(ctx, idents, n_idents, &ident_pc);
    if (!rc) rc = lsg_sort_spans(ctx, spans, n_spans, idents, n_idents, sorted->key, &span_perm);
    ctx->stream = main_stream;
    if (rc) return rc;
    LSG_CUDA(ctx, cudaEventRecord(ctx->ev_join, ctx->side));
  }
  // the join gathers its own tiles through the permutation: no sorted copy of the events (LSG_JOIN_SORTED_COPY=1
  // keeps the separate gather pass, for comparison)
  static const bool sorted_copy = getenv("LSG_JOIN_SORTED_COPY") != nullptr;
  LSG_TRY(lsg_sort_events_finish(ctx, events, n_events, idents, n_idents, nullptr, sorted, true, sorted_copy));
  LSG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
  JoinArgs ja;
  ja.spans = spans; ja.n_spans = n_spans; ja.span_perm = span_perm;
  ja.ev = *sorted; ja.events_in = events; ja.idents = idents; ja.n_idents = n_idents;
  ja.window_ns = window_ns; ja.threshold = threshold; ja.fanout = fanout;
  ja.attr_base = attr_base; ja.cand_event = cand_event; ja.cand_tier = cand_tier; ja.attr_out = attr_out;
  ja.conf_out = conf_out; ja.retrieval_out = retrieval_out; ja.n_cand_out = n_cand_out;
  ja.counters = (uint64_t *)((char *)small_v + SMALL_JOIN);
  ja.inc_keys = inc_keys; ja.n_incidents = n_incidents;
  ja.ident_pc = ident_pc;
  ja.cand_value = cand_value; ja.cand_sig = cand_sig;
  ja.feat = feat;
  LSG_TRY(lsg_launch_join(ctx, ja));
  *join_counters_dev = ja.counters;
  return 0;
}

void fill_debug(lsg_debug *d, const uint64_t *jc, const SortedEvents &sorted, uint64_t n_unsupported, size_t n_spans) {
  const uint64_t n_supported = (uint64_t)sorted.n - n_unsupported;
  d->unsupported_type = n_unsupported * (uint64_t)n_spans;   // correlator.go:73-77, once per span
  d->unmatched = n_supported * (uint64_t)n_spans - jc[0];           // every supported pair that did not match
  d->low_confidence = jc[1];
  d->fanout_dropped = jc[2];
}

__global__ void k_finalize_counters(const unsigned long long *__restrict__ jc, unsigned long long n_events,
                                    const LsgBatchStats *__restrict__ stats, unsigned long long n_spans,
                                    unsigned long long n_incidents, unsigned long long *__restrict__ out) {
  // layout = lsg_counters: debug[4] | confusion | events spans incidents candidates kept bad_idents unsupported_events
  const int tail = 4 + (int)(sizeof(lsg_confusion_out) / 8);
  const unsigned long long n_unsupported = stats->n_unsupported;
  out[tail + 5] = (stats->bad != 0 || jc[5] != 0) ? 1ull : 0ull;
  out[0] = (n_events - n_unsupported) * n_spans - jc[0];
  out[1] = jc[1];
  out[2] = jc[2];
  out[3] = n_unsupported * n_spans;
  out[tail + 0] = n_events;
  out[tail + 1] = n_spans;
  out[tail + 2] = n_incidents;
  out[tail + 3] = jc[3];
  out[tail + 4] = jc[4];
  out[tail + 6] = n_unsupported;
}

__global__ void k_widen_keys(const uint32_t *__restrict__ in, size_t n, unsigned long long *__restrict__ out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = (unsigned long long)in[i];
}

}  // namespace

extern "C" {

const char *lsg_version(void) { return "llmslo_gpu 0.1 (sm_100a)"; }

int lsg_abi_sizes(size_t *out, int n) {
  const size_t v[] = {sizeof(lsg_event), sizeof(lsg_span), sizeof(lsg_ident), sizeof(lsg_debug), sizeof(lsg_eval),
                      sizeof(lsg_model), sizeof(lsg_confusion_out), sizeof(lsg_feature), sizeof(lsg_counters),
                      sizeof(lsg_pipeline_args), sizeof(lsg_synth_cfg), sizeof(lsg_event24)};
  const int m = (int)(sizeof v / sizeof v[0]);
  for (int i = 0; i < n && i < m; ++i) out[i] = v[i];
  return m;
}

int lsg_create(int device, void *stream, lsg_ctx **out) {
  if (!out) return LSG_ERR_INVALID;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0 || device < 0 || device >= count) {
    cudaGetLastError();
    return LSG_ERR_NODEVICE;
  }
  lsg_ctx *ctx = new (std::nothrow) lsg_ctx();
  if (!ctx) return LSG_ERR_NOMEM;
  ctx->device = device;
  auto fail = [&](int code) {   // lsg_destroy releases whatever exists so far
    cudaGetLastError();
    lsg_destroy(ctx);
    return code;
  };
  if (cudaSetDevice(device) != cudaSuccess) return fail(LSG_ERR_NODEVICE);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return fail(LSG_ERR_NODEVICE);
  if (prop.major < 10) return fail(LSG_ERR_NODEVICE);   // the kernels are built for sm_100a only
  ctx->sm_count = prop.multiProcessorCount;
  if (stream) {
    ctx->stream = (cudaStream_t)stream;
  } else {
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) return fail(LSG_ERR_CUDA);
    ctx->own_stream = true;
  }
  if (cudaStreamCreateWithFlags(&ctx->side, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming) != cudaSuccess)
    return fail(LSG_ERR_CUDA);
  ctx->pinned_cap = 8192;
  if (cudaMallocHost(&ctx->pinned, ctx->pinned_cap) != cudaSuccess) return fail(LSG_ERR_NOMEM);
  lsg_model m;
  lsg_default_model(&m);
  build_dev_model(&m, &ctx->model);
  ctx->model_set = true;
  ctx->model_table_valid = false;
  *out = ctx;
  return 0;
}

void lsg_destroy(lsg_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (int i = 0; i < WS_COUNT; ++i)
    if (ctx->ws[i]) cudaFree(ctx->ws[i]);
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  for (auto &sp : ctx->prof_open) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
  for (auto e : ctx->prof_pool) cudaEventDestroy(e);
  if (ctx->side) { cudaStreamSynchronize(ctx->side); cudaStreamDestroy(ctx->side); }
  if (ctx->up) { cudaStreamSynchronize(ctx->up); cudaStreamDestroy(ctx->up); }
  if (ctx->down) { cudaStreamSynchronize(ctx->down); cudaStreamDestroy(ctx->down); }
  for (int s = 0; s < 2; ++s) {
    if (ctx->ev_up[s]) cudaEventDestroy(ctx->ev_up[s]);
    if (ctx->ev_done[s]) cudaEventDestroy(ctx->ev_done[s]);
    if (ctx->ev_down[s]) cudaEventDestroy(ctx->ev_down[s]);
    if (ctx->slot_counters_host[s]) cudaFreeHost(ctx->slot_counters_host[s]);
  }
  if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
  if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

// ---- per-kernel timing ------------------------------------------------------------------------------
static const char *kProfNames[PK_COUNT] = {"event_stats", "event_keys", "radix_hist", "radix_scan", "radix_scatter",
                                           "gather_events", "span_sort", "join", "incident_features", "attribute",
                                           "confusion", "match_pairs", "span_ranges", "bucket_count", "bucket_scatter",
                                           "bucket_spans", "tile_list", "widen_events24"};
int lsg_profile_enable(lsg_ctx *ctx, int on) {
  if (!ctx) return LSG_ERR_INVALID;
  ctx->prof = on != 0;
  return 0;
}
const char *lsg_profile_name(int kind) { return kind >= 0 && kind < PK_COUNT ? kProfNames[kind] : nullptr; }
int lsg_profile_read(lsg_ctx *ctx, double *ms, uint64_t *count, int cap, int reset) {
  if (!ctx) return LSG_ERR_INVALID;
  LSG_TRY(set_device(ctx));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (auto &sp : ctx->prof_open) {
    float t = 0.f;
    if (cudaEventElapsedTime(&t, sp.a, sp.b) == cudaSuccess) {
      ctx->prof_ms[sp.kind] += t;
      ctx->prof_n[sp.kind]++;
    }
    ctx->prof_pool.push_back(sp.a);
    ctx->prof_pool.push_back(sp.b);
  }
  ctx->prof_open.clear();
  for (int i = 0; i < cap && i < PK_COUNT; ++i) {
    if (ms) ms[i] = ctx->prof_ms[i];
    if (count) count[i] = ctx->prof_n[i];
  }
  if (reset)
    for (int i = 0; i < PK_COUNT; ++i) { ctx->prof_ms[i] = 0; ctx->prof_n[i] = 0; }
  return PK_COUNT;
}

int lsg_last_sort_info(const lsg_ctx *ctx, uint32_t *key_bits, uint32_t *time_shift, uint32_t *passes) {
  if (!ctx) return LSG_ERR_INVALID;
  if (key_bits) *key_bits = ctx->last_key.bits;
  if (time_shift) *time_shift = ctx->last_key.shift;
  if (passes) {
    // as radix_sort_pairs shares the key's bits: as few passes as 10-bit digits allow, digits of equal width, never
    // narrower than 8 bits (what the radix path runs; the counting sort of small key spaces has no passes)
    const uint32_t bits = ctx->last_key.bits ? ctx->last_key.bits : 1;
    const uint32_t most = (bits + 9) / 10;
    uint32_t width = (bits + most - 1) / most;
    if (width < 8) width = 8;
    *passes = (bits + width - 1) / width;
  }
  return 0;
}

const char *lsg_last_error(const lsg_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }
int lsg_set_join_path(lsg_ctx *ctx, int path) {
  if (!ctx || path < -1 || path > 3) return LSG_ERR_INVALID;
  ctx->join_path = path;
  return 0;
}
uint64_t lsg_launch_count(const lsg_ctx *ctx) { return ctx ? ctx->launches : 0; }
int lsg_host_alloc(size_t bytes, int write_combined, void **out) {
  if (!out) return LSG_ERR_INVALID;
  *out = nullptr;
  const cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, write_combined ? cudaHostAllocWriteCombined : cudaHostAllocDefault);
  if (e != cudaSuccess) {
    cudaGetLastError();
    *out = nullptr;
    return e == cudaErrorMemoryAllocation ? LSG_ERR_NOMEM : LSG_ERR_CUDA;
  }
  return 0;
}

void lsg_host_free(void *p) {
  if (p) cudaFreeHost(p);
}

size_t lsg_workspace_bytes(const lsg_ctx *ctx) {
  size_t t = 0;
  if (ctx) for (int i = 0; i < WS_COUNT; ++i) t += ctx->ws_cap[i];
  return t;
}

void lsg_default_model(lsg_model *m) {
  // DefaultPriors / DefaultLikelihoods / signalThresholds, bayesian.go:54-207
  static const double lik[LSG_NSIG][LSG_NDOM] = {
      {0.95, 0.70, 0.10, 0.10, 0.10, 0.10, 0.15, 0.10},   // dns_latency_ms
      {0.15, 0.90, 0.10, 0.10, 0.10, 0.15, 0.10, 0.10},   // tcp_retransmits_total
      {0.10, 0.10, 0.90, 0.60, 0.10, 0.10, 0.10, 0.10},   // runqueue_delay_ms
      {0.50, 0.85, 0.10, 0.10, 0.75, 0.40, 0.30, 0.10},   // connect_latency_ms
      {0.10, 0.30, 0.10, 0.10, 0.80, 0.50, 0.20, 0.10},   // tls_handshake_ms
      {0.10, 0.10, 0.90, 0.20, 0.10, 0.10, 0.10, 0.10},   // cpu_steal_pct
      {0.10, 0.10, 0.85, 0.75, 0.10, 0.10, 0.10, 0.10},   // cfs_throttled_ms
      {0.05, 0.05, 0.15, 0.95, 0.05, 0.05, 0.05, 0.05},   // mem_reclaim_latency_ms
      {0.05, 0.05, 0.10, 0.85, 0.05, 0.05, 0.30, 0.05},   // disk_io_latency_ms
      {0.10, 0.20, 0.15, 0.10, 0.90, 0.60, 0.40, 0.10},   // syscall_latency_ms
      {0.10, 0.80, 0.05, 0.05, 0.60, 0.85, 0.15, 0.10},   // connect_errors_total
      {0.05, 0.70, 0.05, 0.05, 0.30, 0.60, 0.10, 0.05}};  // tls_handshake_fail_total
  static const double thr[LSG_NSIG] = {40, 2, 10, 80, 60, 2, 40, 5, 10, 50, 1, 1};
  memset(m, 0, sizeof *m);
  for (int d = 0; d < LSG_NDOM; ++d) m->prior[d] = 1.0 / (double)LSG_NDOM;
  memcpy(m->lik, lik, sizeof lik);
  memcpy(m->thresh, thr, sizeof thr);
  m->sig_mask = (1u << LSG_NSIG) - 1;
}

int lsg_set_model(lsg_ctx *ctx, const lsg_model *model) {
  if (!ctx || !model) return LSG_ERR_INVALID;
  build_dev_model(model, &ctx->model);
  ctx->model_set = true;
  ctx->model_table_valid = false;
  return 0;
}

// ---- A2-A3 ----------------------------------------------------------------------------------------
int lsg_match_pairs_dev(lsg_ctx *ctx, const lsg_span *spans, const lsg_event *events, size_t n, const lsg_ident *idents,
                        size_t n_idents, int64_t window_ns, double threshold, const uint8_t *expected_match,
                        const uint8_t *expected_tier, uint8_t *tier_out, uint8_t *predicted_out, lsg_eval *eval_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!spans || !events || !idents || !tier_out || !predicted_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  LSG_TRY(norm_params(ctx, &window_ns, &threshold, nullptr));
  void *small_v;
  LSG_TRY(lsg_ws(ctx, WS_SMALL, 4096, &small_v));
  uint64_t *eval_dev = (uint64_t *)((char *)small_v + SMALL_EVAL);
  LSG_TRY(lsg_launch_match_pairs(ctx, spans, events, n, idents, n_idents, window_ns, threshold, expected_match, expected_tier,
                                 tier_out, predicted_out, eval_dev));
  uint64_t *pin = (uint64_t *)ctx->pinned;
  const size_t words = sizeof(lsg_eval) / 8 + 1;
  LSG_CUDA(ctx, cudaMemcpyAsync(pin, eval_dev, words * 8, cudaMemcpyDeviceToHost, ctx->stream));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (pin[words - 1] & 0xffffffffu) return lsg_fail(ctx, LSG_ERR_RANGE, "a record's ident is outside the identity table (n_idents=%zu)", n_idents);
  if (eval_out) memcpy(eval_out, pin, sizeof(lsg_eval));
  return 0;
}

int lsg_match_pairs(lsg_ctx *ctx, const lsg_span *spans, const lsg_event *events, size_t n, const lsg_ident *idents,
                    size_t n_idents, int64_t window_ns, double threshold, const uint8_t *expected_match,
                    const uint8_t *expected_tier, uint8_t *tier_out, uint8_t *predicted_out, lsg_eval *eval_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!spans || !events || !idents || !tier_out || !predicted_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  const lsg_span *d_sp; const lsg_event *d_ev; const lsg_ident *d_id; const uint8_t *d_em, *d_et;
  LSG_TRY(h2d(ctx, WS_SPANS_IN, spans, n, &d_sp));
  LSG_TRY(h2d(ctx, WS_EVENTS_IN, events, n, &d_ev));
  LSG_TRY(h2d(ctx, WS_IDENTS, idents, n_idents, &d_id));
  LSG_TRY(h2d(ctx, WS_IN0, expected_match, n, &d_em));
  LSG_TRY(h2d(ctx, WS_IN1, expected_tier, n, &d_et));
  uint8_t *d_tier, *d_pred;
  LSG_TRY(dalloc(ctx, WS_OUT0, true, n, &d_tier));
  LSG_TRY(dalloc(ctx, WS_OUT1, true, n, &d_pred));
  // kernel, then results; lsg_match_pairs_dev synchronises after reading the counters back
  int rc = lsg_match_pairs_dev(ctx, d_sp, d_ev, n, d_id, n_idents, window_ns, threshold, d_em, d_et, d_tier, d_pred, eval_out);
  if (rc) return rc;
  LSG_TRY(d2h(ctx, tier_out, (const uint8_t *)d_tier, n));
  LSG_TRY(d2h(ctx, predicted_out, (const uint8_t *)d_pred, n));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- A4-A6 ----------------------------------------------------------------------------------------
int lsg_process_batch_dev(lsg_ctx *ctx, const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                          const lsg_ident *idents, size_t n_idents, int64_t window_ns, double threshold, int fanout,
                          const double *attr_base, uint32_t *cand_event, uint8_t *cand_tier, double *attr_out, double *conf_out,
                          double *retrieval_out, uint32_t *n_cand_out, lsg_debug *debug_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if ((n_spans && (!spans || !cand_event || !cand_tier || !attr_out || !conf_out)) || (n_events && !events) ||
      ((n_spans || n_events) && !idents))
    return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  LSG_TRY(norm_params(ctx, &window_ns, &threshold, &fanout));
  uint64_t *jc_dev = nullptr;
  SortedEvents sorted;
  LSG_TRY(process_batch_core(ctx, spans, n_spans, events, n_events, idents, n_idents, window_ns, threshold, fanout, attr_base,
                             cand_event, cand_tier, attr_out, conf_out, retrieval_out, n_cand_out, &jc_dev, &sorted));
  // one readback: the batch statistics (start of WS_SMALL) and the join's counters behind them
  char *pin = (char *)ctx->pinned;
  LSG_CUDA(ctx, cudaMemcpyAsync(pin, sorted.stats_dev, SMALL_JOIN + 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const LsgBatchStats st = *(const LsgBatchStats *)pin;
  const uint64_t *jc = (const uint64_t *)(pin + SMALL_JOIN);
  if (st.bad & 1u) return lsg_fail(ctx, LSG_ERR_RANGE, "an event's ident is outside the identity table (n_idents=%zu)", n_idents);
  if ((st.bad & 2u) || jc[5]) return lsg_fail(ctx, LSG_ERR_RANGE, "a span's ident is outside the identity table (n_idents=%zu)", n_idents);
  if (debug_out) fill_debug(debug_out, jc, sorted, st.n_unsupported, n_spans);
  return 0;
}

int lsg_process_batch(lsg_ctx *ctx, const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                      const lsg_ident *idents, size_t n_idents, int64_t window_ns, double threshold, int fanout,
                      const double *attr_base, uint32_t *cand_event, uint8_t *cand_tier, double *attr_out, double *conf_out,
                      double *retrieval_out, uint32_t *n_cand_out, lsg_debug *debug_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if ((n_spans && (!spans || !cand_event || !cand_tier || !attr_out || !conf_out)) || (n_events && !events) ||
      ((n_spans || n_events) && !idents))
    return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  int f = fanout <= 0 ? 3 : fanout;
  const lsg_span *d_sp; const lsg_event *d_ev; const lsg_ident *d_id; const double *d_base;
  LSG_TRY(h2d(ctx, WS_SPANS_IN, spans, n_spans, &d_sp));
  LSG_TRY(h2d(ctx, WS_EVENTS_IN, events, n_events, &d_ev));
  LSG_TRY(h2d(ctx, WS_IDENTS, idents, n_idents, &d_id));
  LSG_TRY(h2d(ctx, WS_IN0, attr_base, n_spans * LSG_NSIG, &d_base));
  uint32_t *d_ce, *d_nc; uint8_t *d_ct; double *d_attr, *d_conf, *d_ret;
  LSG_TRY(dalloc(ctx, WS_OUT0, true, n_spans * f, &d_ce));
  LSG_TRY(dalloc(ctx, WS_OUT1, true, n_spans * f, &d_ct));
  LSG_TRY(dalloc(ctx, WS_OUT2, true, n_spans * LSG_NSIG, &d_attr));
  LSG_TRY(dalloc(ctx, WS_OUT3, true, n_spans, &d_conf));
  LSG_TRY(dalloc(ctx, WS_OUT4, retrieval_out != nullptr, n_spans, &d_ret));
  LSG_TRY(dalloc(ctx, WS_OUT5, n_cand_out != nullptr, n_spans, &d_nc));
  int rc = lsg_process_batch_dev(ctx, d_sp, n_spans, d_ev, n_events, d_id, n_idents, window_ns, threshold, fanout, d_base, d_ce,
                                 d_ct, d_attr, d_conf, d_ret, d_nc, debug_out);
  if (rc) return rc;
  LSG_TRY(d2h(ctx, cand_event, (const uint32_t *)d_ce, n_spans * f));
  LSG_TRY(d2h(ctx, cand_tier, (const uint8_t *)d_ct, n_spans * f));
  LSG_TRY(d2h(ctx, attr_out, (const double *)d_attr, n_spans * LSG_NSIG));
  LSG_TRY(d2h(ctx, conf_out, (const double *)d_conf, n_spans));
  LSG_TRY(d2h(ctx, retrieval_out, (const double *)d_ret, n_spans));
  LSG_TRY(d2h(ctx, n_cand_out, (const uint32_t *)d_nc, n_spans));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

int lsg_sort_events_dev(lsg_ctx *ctx, const lsg_event *events, size_t n_events, const lsg_ident *idents, size_t n_idents,
                        lsg_event *sorted_out, uint32_t *perm_out, uint64_t *keys_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n_events && (!events || !idents || !sorted_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  SortedEvents sorted;
  LSG_TRY(lsg_sort_events(ctx, events, n_events, idents, n_idents, 0, true, sorted_out, &sorted));
  if (n_events && perm_out)
    LSG_CUDA(ctx, cudaMemcpyAsync(perm_out, sorted.perm, n_events * 4, cudaMemcpyDeviceToDevice, ctx->stream));
  if (n_events && keys_out) {
    if (sorted.key.bits <= 31) {   // short keys are sorted as 32-bit words; the ABI always hands out 64-bit keys
      k_widen_keys<<<(unsigned)((n_events + 255) / 256 > 4096 ? 4096 : (n_events + 255) / 256), 256, 0, ctx->stream>>>(
          (const uint32_t *)sorted.keys, n_events, (unsigned long long *)keys_out);
      LSG_LAUNCH_CHECK(ctx);
    } else {
      LSG_CUDA(ctx, cudaMemcpyAsync(keys_out, sorted.keys, n_events * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    }
  }
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- A7 -------------------------------------------------------------------------------------------
int lsg_incident_features_dev(lsg_ctx *ctx, const double *span_attr, const uint32_t *span_incident, size_t n_spans,
                              size_t n_incidents, double *signals_out, lsg_feature *features_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if ((n_spans && (!span_attr || !span_incident)) || (n_incidents && !signals_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  LSG_TRY(lsg_launch_incident_features(ctx, span_attr, span_incident, nullptr, n_spans, n_incidents, signals_out, features_out));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

int lsg_incident_features(lsg_ctx *ctx, const double *span_attr, const uint32_t *span_incident, size_t n_spans,
                          size_t n_incidents, double *signals_out, lsg_feature *features_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if ((n_spans && (!span_attr || !span_incident)) || (n_incidents && !signals_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  const double *d_attr; const uint32_t *d_inc;
  LSG_TRY(h2d(ctx, WS_IN0, span_attr, n_spans * LSG_NSIG, &d_attr));
  LSG_TRY(h2d(ctx, WS_IN1, span_incident, n_spans, &d_inc));
  double *d_sig; lsg_feature *d_feat;
  LSG_TRY(dalloc(ctx, WS_OUT0, true, n_incidents * LSG_NSIG, &d_sig));
  LSG_TRY(dalloc(ctx, WS_OUT1, features_out != nullptr, n_incidents * LSG_NSIG, &d_feat));
  LSG_TRY(lsg_launch_incident_features(ctx, d_attr, d_inc, nullptr, n_spans, n_incidents, d_sig, d_feat));
  LSG_TRY(d2h(ctx, signals_out, (const double *)d_sig, n_incidents * LSG_NSIG));
  LSG_TRY(d2h(ctx, features_out, (const lsg_feature *)d_feat, n_incidents * LSG_NSIG));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- A9-A10 ---------------------------------------------------------------------------------------
int lsg_attribute_dev(lsg_ctx *ctx, const double *signals, const uint8_t *has_signals, const uint8_t *rule_label, size_t n,
                      double *posterior, uint8_t *top1, uint16_t *elev_mask) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!signals || !posterior || !top1)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  LSG_TRY(lsg_launch_attribute(ctx, signals, has_signals, rule_label, n, posterior, top1, elev_mask));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

int lsg_attribute(lsg_ctx *ctx, const double *signals, const uint8_t *has_signals, const uint8_t *rule_label, size_t n,
                  double *posterior, uint8_t *top1, uint16_t *elev_mask) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!signals || !posterior || !top1)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  const double *d_sig; const uint8_t *d_has, *d_rule;
  LSG_TRY(h2d(ctx, WS_IN0, signals, n * LSG_NSIG, &d_sig));
  LSG_TRY(h2d(ctx, WS_IN1, has_signals, n, &d_has));
  LSG_TRY(h2d(ctx, WS_IN2, rule_label, n, &d_rule));
  double *d_post; uint8_t *d_top; uint16_t *d_elev;
  LSG_TRY(dalloc(ctx, WS_OUT0, true, n * LSG_NDOM, &d_post));
  LSG_TRY(dalloc(ctx, WS_OUT1, true, n, &d_top));
  LSG_TRY(dalloc(ctx, WS_OUT2, elev_mask != nullptr, n, &d_elev));
  LSG_TRY(lsg_launch_attribute(ctx, d_sig, d_has, d_rule, n, d_post, d_top, d_elev));
  LSG_TRY(d2h(ctx, posterior, (const double *)d_post, n * LSG_NDOM));
  LSG_TRY(d2h(ctx, top1, (const uint8_t *)d_top, n));
  LSG_TRY(d2h(ctx, elev_mask, (const uint16_t *)d_elev, n));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- A12-A13 --------------------------------------------------------------------------------------
int lsg_confusion_dev(lsg_ctx *ctx, const uint8_t *actual, const uint8_t *predicted, const uint32_t *expected,
                      const uint8_t *expected_n, const double *posterior, double coverage_thr, double hyp_floor, size_t n,
                      lsg_confusion_out *out) {
  if (!ctx || !out) return LSG_ERR_INVALID;
  if (n && (!actual || !predicted || !expected || !expected_n)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  void *small_v;
  LSG_TRY(lsg_ws(ctx, WS_SMALL, 4096, &small_v));
  uint64_t *conf_dev = (uint64_t *)((char *)small_v + SMALL_CONF);
  LSG_CUDA(ctx, cudaMemsetAsync(conf_dev, 0, sizeof(lsg_confusion_out), ctx->stream));
  LSG_TRY(lsg_launch_confusion(ctx, actual, predicted, expected, expected_n, posterior, coverage_thr, hyp_floor, n, conf_dev));
  LSG_CUDA(ctx, cudaMemcpyAsync(ctx->pinned, conf_dev, sizeof(lsg_confusion_out), cudaMemcpyDeviceToHost, ctx->stream));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(out, ctx->pinned, sizeof(lsg_confusion_out));
  return 0;
}

int lsg_confusion(lsg_ctx *ctx, const uint8_t *actual, const uint8_t *predicted, const uint32_t *expected,
                  const uint8_t *expected_n, const double *posterior, double coverage_thr, double hyp_floor, size_t n,
                  lsg_confusion_out *out) {
  if (!ctx || !out) return LSG_ERR_INVALID;
  if (n && (!actual || !predicted || !expected || !expected_n)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  const uint8_t *d_a, *d_p, *d_en; const uint32_t *d_e; const double *d_post;
  LSG_TRY(h2d(ctx, WS_IN0, actual, n, &d_a));
  LSG_TRY(h2d(ctx, WS_IN1, predicted, n, &d_p));
  LSG_TRY(h2d(ctx, WS_IN2, expected, n, &d_e));
  LSG_TRY(h2d(ctx, WS_IN3, expected_n, n, &d_en));
  LSG_TRY(h2d(ctx, WS_IN4, posterior, n * LSG_NDOM, &d_post));
  return lsg_confusion_dev(ctx, d_a, d_p, d_e, d_en, d_post, coverage_thr, hyp_floor, n, out);
}

// ---- the whole hot path, device resident ------------------------------------------------------------
int lsg_pipeline_dev(lsg_ctx *ctx, const lsg_pipeline_args *a) {
  if (!ctx || !a) return LSG_ERR_INVALID;
  // attr_out, conf_out, cand_value and cand_sig are optional outputs
  // an empty array may be NULL (a Go caller's empty slice): ProcessBatch with no signals returns the spans unchanged
  if ((a->n_spans && (!a->spans || !a->cand_event || !a->cand_tier)) || (a->n_events && !a->events && !a->events24) ||
      ((a->n_spans || a->n_events) && !a->idents) || !a->counters_dev ||
      (a->n_incidents && (!a->inc_signals || !a->posterior || !a->top1 || !a->inc_actual || !a->inc_expected ||
                          !a->inc_expected_n)))
    return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  int64_t window_ns = a->window_ns;
  double threshold = a->threshold;
  int fanout = a->fanout;
  LSG_TRY(norm_params(ctx, &window_ns, &threshold, &fanout));
  LSG_CUDA(ctx, cudaMemsetAsync(a->counters_dev, 0, sizeof(lsg_counters), ctx->stream));
  uint64_t *jc_dev = nullptr;
  SortedEvents sorted;
  LsgKeyHint hint{a->span_ts_lo, a->span_ts_hi, a->max_grp_hint};
  // the per-incident maxima (A7) are accumulated by the join's epilogue: no second pass over the span attributes
  void *acc = nullptr;
  LSG_TRY(lsg_ws(ctx, WS_FEAT, a->n_incidents * LSG_NSIG * 8, &acc));
  LSG_CUDA(ctx, cudaMemsetAsync(acc, 0, a->n_incidents * LSG_NSIG * 8, ctx->stream));
  const lsg_event *events = a->events;
  if (a->events24 && a->n_events) {   // the 24-byte upload form: widened once, every kernel reads 32-byte records
    void *wide = nullptr;
    LSG_TRY(lsg_ws(ctx, WS_EV_WIDE, a->n_events * sizeof(lsg_event), &wide));
    LSG_TRY(lsg_launch_widen_events24(ctx, a->events24, a->n_events, a->events24_base, (lsg_event *)wide));
    events = (const lsg_event *)wide;
  }
  if (a->features_out && a->n_incidents)
    LSG_CUDA(ctx, cudaMemsetAsync(a->features_out, 0, a->n_incidents * LSG_NSIG * sizeof(lsg_feature), ctx->stream));
  LSG_TRY(process_batch_core(ctx, a->spans, a->n_spans, events, a->n_events, a->idents, a->n_idents, window_ns, threshold,
                             fanout, nullptr, a->cand_event, a->cand_tier, a->attr_out, a->conf_out, nullptr, nullptr, &jc_dev,
                             &sorted, (unsigned long long *)acc, a->n_incidents, a->cand_value, a->cand_sig,
                             a->key_hint ? &hint : nullptr, a->features_out, a->events24 ? &a->events24_base : nullptr));
  LSG_TRY(lsg_launch_incident_finish(ctx, (const unsigned long long *)acc, a->n_incidents, a->inc_signals));
  LSG_TRY(lsg_launch_attribute(ctx, a->inc_signals, nullptr, nullptr, a->n_incidents, a->posterior, a->top1, nullptr));
  LSG_TRY(lsg_launch_confusion(ctx, a->inc_actual, a->top1, a->inc_expected, a->inc_expected_n, a->posterior, a->coverage_thr,
                               0.01, a->n_incidents, a->counters_dev + 4));
  k_finalize_counters<<<1, 1, 0, ctx->stream>>>((const unsigned long long *)jc_dev, a->n_events, sorted.stats_dev, a->n_spans,
                                                a->n_incidents, (unsigned long long *)a->counters_dev);
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}

// host-buffer form of the whole path (bench.py's e2e): every pointer of `a` is a HOST pointer (pinned memory
// makes the copies asynchronous); inputs go up, the pipeline runs, results come back, then it synchronises.
int lsg_pipeline(lsg_ctx *ctx, const lsg_pipeline_args *a, lsg_counters *counters_out) {
  if (!ctx || !a) return LSG_ERR_INVALID;
  if ((a->n_spans && !a->spans) || (a->n_events && !a->events && !a->events24) || ((a->n_spans || a->n_events) && !a->idents) ||
      (a->n_incidents && (!a->inc_actual || !a->inc_expected || !a->inc_expected_n)))
    return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_TRY(set_device(ctx));
  const int f = a->fanout <= 0 ? 3 : a->fanout;
  lsg_pipeline_args d = *a;
  LSG_TRY(h2d(ctx, WS_SPANS_IN, a->spans, a->n_spans, &d.spans));
  if (a->events24) {
    d.events = nullptr;
    LSG_TRY(h2d(ctx, WS_EVENTS_IN, a->events24, a->n_events, &d.events24));
  } else {
    LSG_TRY(h2d(ctx, WS_EVENTS_IN, a->events, a->n_events, &d.events));
  }
  LSG_TRY(dalloc(ctx, WS_IN5, a->features_out != nullptr, a->n_incidents * LSG_NSIG, &d.features_out));
  LSG_TRY(h2d(ctx, WS_IDENTS, a->idents, a->n_idents, &d.idents));
  LSG_TRY(h2d(ctx, WS_IN0, a->inc_actual, a->n_incidents, &d.inc_actual));
  LSG_TRY(h2d(ctx, WS_IN1, a->inc_expected, a->n_incidents, &d.inc_expected));
  LSG_TRY(h2d(ctx, WS_IN2, a->inc_expected_n, a->n_incidents, &d.inc_expected_n));
  LSG_TRY(dalloc(ctx, WS_OUT0, true, a->n_spans * f, &d.cand_event));
  LSG_TRY(dalloc(ctx, WS_OUT1, true, a->n_spans * f, &d.cand_tier));
  LSG_TRY(dalloc(ctx, WS_OUT2, a->attr_out != nullptr, a->n_spans * LSG_NSIG, &d.attr_out));   // only what the caller takes
  LSG_TRY(dalloc(ctx, WS_OUT3, a->conf_out != nullptr, a->n_spans, &d.conf_out));
  LSG_TRY(dalloc(ctx, WS_OUT6, a->cand_value != nullptr, a->n_spans * f, &d.cand_value));
  LSG_TRY(dalloc(ctx, WS_OUT7, a->cand_sig != nullptr, a->n_spans * f, &d.cand_sig));
  LSG_TRY(dalloc(ctx, WS_OUT4, true, a->n_incidents * LSG_NSIG, &d.inc_signals));
  LSG_TRY(dalloc(ctx, WS_OUT5, true, a->n_incidents * LSG_NDOM, &d.posterior));
  LSG_TRY(dalloc(ctx, WS_IN3, true, a->n_incidents, &d.top1));
  LSG_TRY(dalloc(ctx, WS_IN4, true, (size_t)LSG_COUNTER_WORDS, &d.counters_dev));
  LSG_TRY(lsg_pipeline_dev(ctx, &d));
  LSG_TRY(d2h(ctx, a->cand_event, (const uint32_t *)d.cand_event, a->n_spans * f));
  LSG_TRY(d2h(ctx, a->cand_tier, (const uint8_t *)d.cand_tier, a->n_spans * f));
  LSG_TRY(d2h(ctx, a->attr_out, (const double *)d.attr_out, a->n_spans * LSG_NSIG));
  LSG_TRY(d2h(ctx, a->conf_out, (const double *)d.conf_out, a->n_spans));
  LSG_TRY(d2h(ctx, a->cand_value, (const double *)d.cand_value, a->n_spans * f));
  LSG_TRY(d2h(ctx, a->cand_sig, (const uint8_t *)d.cand_sig, a->n_spans * f));
  LSG_TRY(d2h(ctx, a->inc_signals, (const double *)d.inc_signals, a->n_incidents * LSG_NSIG));
  LSG_TRY(d2h(ctx, a->posterior, (const double *)d.posterior, a->n_incidents * LSG_NDOM));
  LSG_TRY(d2h(ctx, a->top1, (const uint8_t *)d.top1, a->n_incidents));
  LSG_TRY(d2h(ctx, a->features_out, (const lsg_feature *)d.features_out, a->n_incidents * LSG_NSIG));
  LSG_CUDA(ctx, cudaMemcpyAsync(ctx->pinned, d.counters_dev, sizeof(lsg_counters), cudaMemcpyDeviceToHost, ctx->stream));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (counters_out) memcpy(counters_out, ctx->pinned, sizeof(lsg_counters));
  if (a->counters_dev) memcpy(a->counters_dev, ctx->pinned, sizeof(lsg_counters));
  if (((const lsg_counters *)ctx->pinned)->bad_idents)
    return lsg_fail(ctx, LSG_ERR_RANGE, "a record's ident is outside the identity table (n_idents=%zu), or the key hint does not hold",
                    a->n_idents);
  return 0;
}

// ---- double-buffered form of lsg_pipeline -------------------------------------------------------------------
// Two slots.  submit(slot) enqueues the upload of that batch on an upload stream, the kernels on the context's
// stream (after the upload) and the download on a download stream (after the kernels), and returns; wait(slot)
// blocks until the batch's results are in the caller's buffers.  With two batches in flight the upload of one,
// the kernels of the other and the download of the previous one overlap (PCIe is full duplex), so the steady
// state costs max(upload, kernels, download) per batch instead of their sum.  The caller's buffers of a slot
// must stay untouched between submit and wait, and should be pinned (otherwise the copies serialise).
}  // extern "C"
namespace {
enum { PB_EVENTS = 0, PB_SPANS, PB_IDENTS, PB_ACT, PB_EXP, PB_EXPN, PB_CE, PB_CT, PB_ATTR, PB_CONF, PB_SIG, PB_POST, PB_TOP1, PB_CNT, PB_CV, PB_CS,
       PB_FEAT, PB_BUFS };
static_assert(PB_BUFS <= 20, "WS_PIPE reserves 20 buffers per slot");

int pipe_init(lsg_ctx *ctx) {
  if (ctx->up) return 0;
  LSG_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->up, cudaStreamNonBlocking));
  LSG_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->down, cudaStreamNonBlocking));
  for (int s = 0; s < 2; ++s) {
    LSG_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_up[s], cudaEventDisableTiming));
    LSG_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_done[s], cudaEventDisableTiming));
    LSG_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_down[s], cudaEventDisableTiming));
    LSG_CUDA(ctx, cudaMallocHost(&ctx->slot_counters_host[s], sizeof(lsg_counters)));
  }
  return 0;
}
template <typename T>
int pipe_buf(lsg_ctx *ctx, int slot, int which, size_t n, T **out) {
  void *p;
  LSG_TRY(lsg_ws(ctx, WS_PIPE + slot * PB_BUFS + which, n * sizeof(T), &p));
  *out = (T *)p;
  return 0;
}
}  // namespace
extern "C" {

int lsg_pipeline_submit(lsg_ctx *ctx, const lsg_pipeline_args *a, int slot) {
  if (!ctx || !a || slot < 0 || slot > 1) return LSG_ERR_INVALID;
  if ((a->n_spans && !a->spans) || (a->n_events && !a->events && !a->events24) || ((a->n_spans || a->n_events) && !a->idents) ||
      (a->n_incidents && (!a->inc_actual || !a->inc_expected || !a->inc_expected_n)))
    return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  if (ctx->slot_busy[slot]) return lsg_fail(ctx, LSG_ERR_INVALID, "slot %d is still in flight: call lsg_pipeline_wait first", slot);
  LSG_TRY(set_device(ctx));
  LSG_TRY(pipe_init(ctx));
  const int f = a->fanout <= 0 ? 3 : a->fanout;
  lsg_pipeline_args d = *a;
  lsg_span *d_sp; lsg_event *d_ev; lsg_ident *d_id; uint8_t *d_act, *d_expn; uint32_t *d_exp;
  LSG_TRY(pipe_buf(ctx, slot, PB_SPANS, a->n_spans, &d_sp));
  LSG_TRY(pipe_buf(ctx, slot, PB_EVENTS, a->n_events, &d_ev));   // (24-byte records use the front of the same buffer)
  d.features_out = nullptr;
  if (a->features_out) LSG_TRY(pipe_buf(ctx, slot, PB_FEAT, a->n_incidents * LSG_NSIG, &d.features_out));
  LSG_TRY(pipe_buf(ctx, slot, PB_IDENTS, a->n_idents, &d_id));
  LSG_TRY(pipe_buf(ctx, slot, PB_ACT, a->n_incidents, &d_act));
  LSG_TRY(pipe_buf(ctx, slot, PB_EXP, a->n_incidents, &d_exp));
  LSG_TRY(pipe_buf(ctx, slot, PB_EXPN, a->n_incidents, &d_expn));
  LSG_TRY(pipe_buf(ctx, slot, PB_CE, a->n_spans * f, &d.cand_event));
  LSG_TRY(pipe_buf(ctx, slot, PB_CT, a->n_spans * f, &d.cand_tier));
  d.attr_out = nullptr; d.conf_out = nullptr; d.cand_value = nullptr; d.cand_sig = nullptr;   // only what the caller takes
  if (a->attr_out) LSG_TRY(pipe_buf(ctx, slot, PB_ATTR, a->n_spans * LSG_NSIG, &d.attr_out));
  if (a->conf_out) LSG_TRY(pipe_buf(ctx, slot, PB_CONF, a->n_spans, &d.conf_out));
  if (a->cand_value) LSG_TRY(pipe_buf(ctx, slot, PB_CV, a->n_spans * f, &d.cand_value));
  if (a->cand_sig) LSG_TRY(pipe_buf(ctx, slot, PB_CS, a->n_spans * f, &d.cand_sig));
  LSG_TRY(pipe_buf(ctx, slot, PB_SIG, a->n_incidents * LSG_NSIG, &d.inc_signals));
  LSG_TRY(pipe_buf(ctx, slot, PB_POST, a->n_incidents * LSG_NDOM, &d.posterior));
  LSG_TRY(pipe_buf(ctx, slot, PB_TOP1, a->n_incidents, &d.top1));
  LSG_TRY(pipe_buf(ctx, slot, PB_CNT, (size_t)LSG_COUNTER_WORDS, &d.counters_dev));
  // upload
  auto up = [&](void *dst, const void *src, size_t bytes) -> int {
    if (bytes) LSG_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->up));
    return 0;
  };
  if (a->events24) LSG_TRY(up(d_ev, a->events24, a->n_events * sizeof(lsg_event24)));
  else LSG_TRY(up(d_ev, a->events, a->n_events * sizeof(lsg_event)));
  LSG_TRY(up(d_sp, a->spans, a->n_spans * sizeof(lsg_span)));
  LSG_TRY(up(d_id, a->idents, a->n_idents * sizeof(lsg_ident)));
  LSG_TRY(up(d_act, a->inc_actual, a->n_incidents));
  LSG_TRY(up(d_exp, a->inc_expected, a->n_incidents * 4));
  LSG_TRY(up(d_expn, a->inc_expected_n, a->n_incidents));
  LSG_CUDA(ctx, cudaEventRecord(ctx->ev_up[slot], ctx->up));
  // kernels
  d.events24 = a->events24 ? (const lsg_event24 *)d_ev : nullptr;
  d.spans = d_sp; d.events = a->events24 ? nullptr : d_ev; d.idents = d_id; d.inc_actual = d_act; d.inc_expected = d_exp; d.inc_expected_n = d_expn;
  LSG_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_up[slot], 0));
  LSG_TRY(lsg_pipeline_dev(ctx, &d));
  LSG_CUDA(ctx, cudaEventRecord(ctx->ev_done[slot], ctx->stream));
  // download
  LSG_CUDA(ctx, cudaStreamWaitEvent(ctx->down, ctx->ev_done[slot], 0));
  auto down = [&](void *dst, const void *src, size_t bytes) -> int {
    if (dst && bytes) LSG_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->down));
    return 0;
  };
  LSG_TRY(down(a->cand_event, d.cand_event, a->n_spans * f * 4));
  LSG_TRY(down(a->cand_tier, d.cand_tier, a->n_spans * f));
  LSG_TRY(down(a->attr_out, d.attr_out, a->n_spans * LSG_NSIG * 8));
  LSG_TRY(down(a->conf_out, d.conf_out, a->n_spans * 8));
  LSG_TRY(down(a->cand_value, d.cand_value, a->n_spans * f * 8));
  LSG_TRY(down(a->cand_sig, d.cand_sig, a->n_spans * f));
  LSG_TRY(down(a->inc_signals, d.inc_signals, a->n_incidents * LSG_NSIG * 8));
  LSG_TRY(down(a->posterior, d.posterior, a->n_incidents * LSG_NDOM * 8));
  LSG_TRY(down(a->top1, d.top1, a->n_incidents));
  LSG_TRY(down(a->features_out, d.features_out, a->n_incidents * LSG_NSIG * sizeof(lsg_feature)));
  LSG_TRY(down(ctx->slot_counters_host[slot], d.counters_dev, sizeof(lsg_counters)));
  LSG_CUDA(ctx, cudaEventRecord(ctx->ev_down[slot], ctx->down));
  ctx->slot_busy[slot] = true;
  return 0;
}

}  // extern "C"
// multi.cu: the device copy of a slot's counter block (what the multi-GPU entry point allreduces)
void *lsg_internal_slot_counters_dev(lsg_ctx *ctx, int slot) { return ctx->ws[WS_PIPE + slot * PB_BUFS + PB_CNT]; }
extern "C" {

int lsg_pipeline_wait(lsg_ctx *ctx, int slot, lsg_counters *counters_out) {
  if (!ctx || slot < 0 || slot > 1) return LSG_ERR_INVALID;
  if (!ctx->slot_busy[slot]) return lsg_fail(ctx, LSG_ERR_INVALID, "slot %d has nothing in flight", slot);
  LSG_TRY(set_device(ctx));
  LSG_CUDA(ctx, cudaEventSynchronize(ctx->ev_down[slot]));
  ctx->slot_busy[slot] = false;
  if (counters_out) memcpy(counters_out, ctx->slot_counters_host[slot], sizeof(lsg_counters));
  if (((const lsg_counters *)ctx->slot_counters_host[slot])->bad_idents)
    return lsg_fail(ctx, LSG_ERR_RANGE, "a record's ident in the batch of slot %d is outside the identity table, or the key hint does not hold", slot);
  return 0;
}

}  // extern "C"
