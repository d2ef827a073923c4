// attribute.cu — K-features (A7 glue), K-bayes (pkg/attribution/bayesian.go:218-284, 315-343) and
// K-confusion (pkg/attribution/pipeline.go:65-185).
#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <mutex>

#include "common.cuh"

namespace {

// ---- K-features ------------------------------------------------------------------------------------
// Per incident and signal: max over the incident's spans of the spans' enriched attribute (the only
// aggregate the reference defines: correlator.go:109-111), plus count / sum / log2 sketch.
// doubles are mapped to order-preserving unsigned keys (f64_key, common.cuh) so that max is an integer atomicMax.

// one thread per (span, signal) element of the [n_spans][12] attribute matrix: coalesced 8-byte loads.
// Lanes of a warp that hit the same (incident, signal) are combined by shuffles first (segmented
// reduction), so runs of spans of one incident cost one atomic per signal per warp.
__global__ void __launch_bounds__(256)
k_incident_accumulate(const double *__restrict__ span_attr, const uint32_t *__restrict__ span_incident,
                      const lsg_span *__restrict__ spans, size_t n_spans, size_t n_incidents,
                      unsigned long long *__restrict__ max_keys, lsg_feature *__restrict__ feat) {
  const size_t total = n_spans * LSG_NSIG;
  const unsigned lane = threadIdx.x & 31;
  for (size_t base = ((size_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31u)); base < total;
       base += (size_t)gridDim.x * blockDim.x) {
    const size_t t = base + lane;
    unsigned long long slot = ~0ull;
    double v = 0.0;
    if (t < total) {
      const size_t i = t / LSG_NSIG;
      const unsigned s = (unsigned)(t - i * LSG_NSIG);
      const uint32_t inc = span_incident ? span_incident[i] : spans[i].incident;
      v = span_attr[t];
      if (inc != LSG_NONE && inc < n_incidents && !isnan(v)) slot = (unsigned long long)inc * LSG_NSIG + s;
    }
    const unsigned m = __match_any_sync(0xffffffffu, slot);
    if (slot == ~0ull) continue;   // no divergence hazard: the remaining shuffles use the member mask
    // reduce within the member group: max key, count, sum (ascending lane order = span order)
    unsigned long long k = f64_key(v);
    unsigned cnt = __popc(m);
    const unsigned leader = __ffs(m) - 1;
    double sum = 0.0;
    unsigned long long kmax = 0;
    for (unsigned mm = m; mm; mm &= mm - 1) {
      const unsigned src = __ffs(mm) - 1;
      const unsigned long long ok = __shfl_sync(m, k, src);
      const double ov = __shfl_sync(m, v, src);
      kmax = ok > kmax ? ok : kmax;
      sum += ov;
    }
    if (lane == leader) atomicMax(&max_keys[slot], kmax);
    if (feat) {
      if (lane == leader) {
        atomicAdd(&feat[slot].count, cnt);
        atomicAdd(&feat[slot].sum, sum);
      }
      int bin = 0;
      if (v > 0) {
        int ex;
        frexp(v, &ex);
        bin = ex - 1 + 4;
        bin = bin < 0 ? 0 : (bin > LSG_SKETCH_BINS - 1 ? LSG_SKETCH_BINS - 1 : bin);
      }
      atomicAdd(&feat[slot].sketch[bin], 1u);
    }
  }
}

__global__ void __launch_bounds__(256) k_incident_finish(const unsigned long long *__restrict__ max_keys, size_t n,
                                                         double *__restrict__ signals_out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    signals_out[i] = key_f64(max_keys[i]);
}

// ---- K-bayes ---------------------------------------------------------------------------------------
// elevated[s] = value >= thresh[s] (bayesian.go:220-225);
// logP[d] = log prior + sum_s log clamp(elevated ? p : 1-p) (bayesian.go:230-242) with the logs taken
// on the host, summed in the fixed order prior, signals 0..11, extra; log-sum-exp in AllDomains order
// (bayesian.go:245-256); top-1 = first maximum of the posteriors (sort.SliceStable, bayesian.go:279-281).
//
// Given the model, the posteriors are a function of the 12-bit elevated mask alone: 4096 possible rows.  Small batches
// compute them per incident (k_attribute); large ones (attribution alone over tens of millions of incidents, where the
// fp64 exp / log of the direct form is what bounds the kernel, not HBM) read them from a table that the SAME device
// function filled for every mask (k_attribute_table), so both forms return identical bits.
__device__ __forceinline__ int posterior_of_mask(const DevModel &m, const unsigned elev, double po[LSG_NDOM]) {
  double lp[LSG_NDOM];
  double mx = -INFINITY;
#pragma unroll
  for (int d = 0; d < LSG_NDOM; ++d) {
    double a = m.log_prior[d];
#pragma unroll
    for (int s = 0; s < LSG_NSIG; ++s) {
      if (m.sig_mask >> s & 1) a = __dadd_rn(a, (elev >> s & 1) ? m.log_elev[s][d] : m.log_not[s][d]);
    }
    a = __dadd_rn(a, m.log_extra[d]);
    lp[d] = a;
    mx = a > mx ? a : mx;
  }
  double se = 0.0;
#pragma unroll
  for (int d = 0; d < LSG_NDOM; ++d) se = __dadd_rn(se, exp(lp[d] - mx));
  const double logz = mx + log(se);
  int best = 0;
  double bestp = -1.0;
#pragma unroll
  for (int d = 0; d < LSG_NDOM; ++d) {
    const double pd = exp(lp[d] - logz);
    po[d] = pd;
    if (pd > bestp) { bestp = pd; best = d; }
  }
  return best;
}

constexpr unsigned AT_MASKS = 1u << LSG_NSIG;

__global__ void __launch_bounds__(128)
k_attribute_table(const __grid_constant__ DevModel m, double *__restrict__ table, uint8_t *__restrict__ best) {
  const unsigned e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= AT_MASKS) return;
  double po[LSG_NDOM];
  best[e] = (uint8_t)posterior_of_mask(m, e, po);
#pragma unroll
  for (int d = 0; d < LSG_NDOM; ++d) table[(size_t)e * LSG_NDOM + d] = po[d];
}

// One thread per incident.  TABLE: the posteriors come from k_attribute_table's rows (L2-resident, 256 KB).
template <bool TABLE>
__global__ void __launch_bounds__(128)
k_attribute(const __grid_constant__ DevModel m, const double *__restrict__ signals, const uint8_t *__restrict__ has_signals,
            const uint8_t *__restrict__ rule_label, size_t n, double *__restrict__ posterior, uint8_t *__restrict__ top1,
            uint16_t *__restrict__ elev_mask, const double *__restrict__ table, const uint8_t *__restrict__ table_best) {
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const double *sg = signals + i * LSG_NSIG;
    unsigned elev = 0;
    bool any = false;
#pragma unroll
    for (int s = 0; s < LSG_NSIG; ++s) {
      const double v = sg[s];
      any |= !isnan(v);
      if (v >= m.thresh[s]) elev |= 1u << s;   // false for NaN value or NaN threshold
    }
    const bool has = has_signals ? has_signals[i] != 0 : any;
    if (!has) {   // AttributeSample returns the rule envelope unchanged (bayesian.go:318-320)
#pragma unroll
      for (int d = 0; d < LSG_NDOM; ++d) posterior[i * LSG_NDOM + d] = nan;
      top1[i] = rule_label ? rule_label[i] : 0xFF;
      if (elev_mask) elev_mask[i] = 0;
      continue;
    }
    double po[LSG_NDOM];
    int best;
    if (TABLE) {
      const double2 *row = reinterpret_cast<const double2 *>(table + (size_t)elev * LSG_NDOM);
#pragma unroll
      for (int d = 0; d < LSG_NDOM / 2; ++d) {
        const double2 v = __ldg(row + d);
        po[2 * d] = v.x; po[2 * d + 1] = v.y;
      }
      best = __ldg(table_best + elev);
    } else {
      best = posterior_of_mask(m, elev, po);
    }
    double2 *out = reinterpret_cast<double2 *>(posterior + i * LSG_NDOM);
#pragma unroll
    for (int d = 0; d < LSG_NDOM / 2; ++d) out[d] = make_double2(po[2 * d], po[2 * d + 1]);
    top1[i] = (uint8_t)best;
    if (elev_mask) elev_mask[i] = (uint16_t)elev;
  }
}

// ---- K-confusion -----------------------------------------------------------------------------------
// counters: lsg_confusion_out as 64-bit words: matrix[16][8] | n | correct | partial | coverage[9][9]
constexpr int CF_MATRIX = 0;
constexpr int CF_N = LSG_MAX_ACTUAL * LSG_NDOM;
constexpr int CF_CORRECT = CF_N + 1;
constexpr int CF_PARTIAL = CF_N + 2;
constexpr int CF_COVER = CF_N + 3;
constexpr int CF_WORDS = CF_COVER + (LSG_NDOM + 1) * (LSG_NDOM + 1);
static_assert(CF_WORDS * 8 == sizeof(lsg_confusion_out), "layout");

__device__ __forceinline__ void warp_agg_add(unsigned int *s_cnt, int slot, bool valid) {
  // warp-aggregated shared-memory atomic: lanes that hit the same counter elect one leader
  const unsigned m = __match_any_sync(0xffffffffu, valid ? slot : -1);
  if (valid && (threadIdx.x & 31) == (unsigned)(__ffs(m) - 1)) atomicAdd(&s_cnt[slot], __popc(m));
}

__global__ void __launch_bounds__(256)
k_confusion(const uint8_t *__restrict__ actual, const uint8_t *__restrict__ predicted, const uint32_t *__restrict__ expected,
            const uint8_t *__restrict__ expected_n, const double *__restrict__ posterior, double coverage_thr,
            double hyp_floor, size_t n, unsigned long long *__restrict__ out) {
  __shared__ unsigned int s_cnt[CF_WORDS];
  for (int i = threadIdx.x; i < CF_WORDS; i += blockDim.x) s_cnt[i] = 0;
  __syncthreads();
  for (size_t base = (size_t)blockIdx.x * blockDim.x; base < n; base += (size_t)gridDim.x * blockDim.x) {
    const size_t i = base + threadIdx.x;
    const bool valid = i < n;
    unsigned a = 0, p = 0, found = 0, en = 0;
    bool part = false;
    if (valid) {
      a = actual[i];
      p = predicted[i];
      unsigned hyp = 0;
      if (posterior) {
#pragma unroll
        for (int d = 0; d < LSG_NDOM; ++d) {
          const double q = posterior[i * LSG_NDOM + d];
          // a hypothesis exists when the posterior is >= hyp_floor (bayesian.go:326-328) and counts for
          // coverage when >= coverage_thr (pipeline.go:163-167); NaN fails both
          if (!(q < hyp_floor) && q >= coverage_thr) hyp |= 1u << d;
        }
      }
      if (p < LSG_NDOM) hyp |= 1u << p;             // pipeline.go:168
      en = expected_n[i];
      const uint32_t ex = expected[i];
      for (unsigned k = 0; k < en && k < 8; ++k) {
        const unsigned dc = ex >> (4 * k) & 15u;
        if (dc < LSG_NDOM) {
          part |= dc == p;                          // pipeline.go:127-132
          found += hyp >> dc & 1u;                  // pipeline.go:170-175
        }
      }
    }
    warp_agg_add(s_cnt, CF_MATRIX + (int)(a * LSG_NDOM + p), valid && a < LSG_MAX_ACTUAL && p < LSG_NDOM);
    warp_agg_add(s_cnt, CF_N, valid);
    warp_agg_add(s_cnt, CF_CORRECT, valid && a == p);
    warp_agg_add(s_cnt, CF_PARTIAL, valid && part);
    warp_agg_add(s_cnt, CF_COVER + (int)(found * (LSG_NDOM + 1) + en), valid && en <= LSG_NDOM);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < CF_WORDS; i += blockDim.x)
    if (s_cnt[i]) atomicAdd(&out[i], (unsigned long long)s_cnt[i]);
}

inline unsigned grid_for(lsg_ctx *ctx, size_t n, int block, int per_sm) {
  size_t blocks = (n + block - 1) / block;
  size_t cap = (size_t)ctx->sm_count * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  return (unsigned)blocks;
}

}  // namespace

int lsg_launch_incident_finish(lsg_ctx *ctx, const unsigned long long *keys, size_t n_incidents, double *signals_out) {
  const size_t n = n_incidents * LSG_NSIG;
  if (n == 0) return 0;
  ProfScope ps(ctx, PK_FEATURES);
  k_incident_finish<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(keys, n, signals_out);
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}

int lsg_launch_incident_features(lsg_ctx *ctx, const double *span_attr, const uint32_t *span_incident, const lsg_span *spans,
                                 size_t n_spans, size_t n_incidents, double *signals_out, lsg_feature *features_out) {
  if (n_incidents == 0) return 0;
  const size_t n = n_incidents * LSG_NSIG;
  void *acc = nullptr;
  LSG_TRY(lsg_ws(ctx, WS_FEAT, n * 8, &acc));
  ProfScope ps(ctx, PK_FEATURES);
  LSG_CUDA(ctx, cudaMemsetAsync(acc, 0, n * 8, ctx->stream));
  if (features_out) LSG_CUDA(ctx, cudaMemsetAsync(features_out, 0, n * sizeof(lsg_feature), ctx->stream));
  if (n_spans) {
    k_incident_accumulate<<<grid_for(ctx, n_spans * LSG_NSIG, 256, 8), 256, 0, ctx->stream>>>(
        span_attr, span_incident, spans, n_spans, n_incidents, (unsigned long long *)acc, features_out);
    LSG_LAUNCH_CHECK(ctx);
  }
  k_incident_finish<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>((const unsigned long long *)acc, n, signals_out);
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}

constexpr size_t AT_TABLE_MIN = 1u << 16;   // incidents from which the mask table pays for its 4096-row build

int lsg_launch_attribute(lsg_ctx *ctx, const double *signals, const uint8_t *has_signals, const uint8_t *rule_label, size_t n,
                         double *posterior, uint8_t *top1, uint16_t *elev_mask) {
  if (n == 0) return 0;
  ProfScope ps(ctx, PK_ATTRIBUTE);
  static const bool no_table = getenv("LSG_ATTRIBUTE_DIRECT") != nullptr;   // comparison knob: always the per-incident form
  if (n >= AT_TABLE_MIN && !no_table) {
    void *tv = nullptr;
    LSG_TRY(lsg_ws(ctx, WS_MODEL_TABLE, (size_t)AT_MASKS * (LSG_NDOM * sizeof(double) + 1), &tv));
    double *table = (double *)tv;
    uint8_t *best = (uint8_t *)(table + (size_t)AT_MASKS * LSG_NDOM);
    if (!ctx->model_table_valid) {
      k_attribute_table<<<AT_MASKS / 128, 128, 0, ctx->stream>>>(ctx->model, table, best);
      LSG_LAUNCH_CHECK(ctx);
      ctx->model_table_valid = true;
    }
    k_attribute<true><<<grid_for(ctx, n, 128, 16), 128, 0, ctx->stream>>>(ctx->model, signals, has_signals, rule_label, n, posterior,
                                                                         top1, elev_mask, table, best);
  } else {
    k_attribute<false><<<grid_for(ctx, n, 128, 16), 128, 0, ctx->stream>>>(ctx->model, signals, has_signals, rule_label, n,
                                                                          posterior, top1, elev_mask, nullptr, nullptr);
  }
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}

int lsg_launch_confusion(lsg_ctx *ctx, const uint8_t *actual, const uint8_t *predicted, const uint32_t *expected,
                         const uint8_t *expected_n, const double *posterior, double coverage_thr, double hyp_floor, size_t n,
                         uint64_t *out_dev) {
  if (n == 0) return 0;
  ProfScope ps(ctx, PK_CONFUSION);
  k_confusion<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(actual, predicted, expected, expected_n, posterior,
                                                                  coverage_thr, hyp_floor, n, (unsigned long long *)out_dev);
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}

// ---- per-incident signal reduction without any allocation ---------------------------------------------
//
// lsg_incident_features[_dev] above keep their accumulators in a context workspace that is cudaMalloc'ed
// on first use and whenever a batch outgrows it.  Some hosts cannot do that: a process that has already
// handed the device's whole memory budget to its own pool has nothing left for cudaMalloc to give, and a
// service that keeps other streams busy cannot afford cudaMalloc/cudaFree, which synchronise the whole
// device.  For them the accumulators below live in the library image itself: a fixed array in device
// global memory that belongs to this file's CUDA module (LSG_STATIC_MAX_INCIDENTS x 12 x 8 bytes = 1.5 MiB per
// GPU), so the entry point never allocates.  With CUDA's default lazy module loading the driver places the module,
// array included, on a GPU at the first symbol lookup or kernel launch from this file on that GPU and keeps it
// until the process exits: that first use is the one moment device memory is taken, so it belongs before a
// caller gives the rest of the device to its own pool (or run with CUDA_MODULE_LOADING=EAGER).
static __device__ __align__(256) unsigned long long g_static_incident_keys[(size_t)LSG_STATIC_MAX_INCIDENTS * LSG_NSIG];

namespace {
// In host code `g_static_incident_keys` is only a handle for the runtime, not an address a kernel could
// use; the device address differs per GPU and is looked up once per device with that device current.
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
unsigned long long *g_static_keys_dev[kStaticMaxDevices] = {};
int g_static_sm_count[kStaticMaxDevices] = {};

void static_err(char *buf, size_t len, const char *fmt, ...) {
  if (!buf || len == 0) return;
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, len, fmt, ap);
  va_end(ap);
}
}  // namespace

extern "C" int lsg_incident_signals_static(void *stream, const double *span_attr, const uint32_t *span_incident,
                                           size_t n_spans, size_t n_incidents, double *signals_out, char *errbuf,
                                           size_t errbuf_len) {
  if (errbuf && errbuf_len) errbuf[0] = 0;
  if ((n_spans && (!span_attr || !span_incident)) || (n_incidents && !signals_out)) {
    static_err(errbuf, errbuf_len, "null argument");
    return LSG_ERR_INVALID;
  }
  if (n_incidents > (size_t)LSG_STATIC_MAX_INCIDENTS) {
    static_err(errbuf, errbuf_len, "n_incidents %zu exceeds LSG_STATIC_MAX_INCIDENTS (%d): the built-in scratch array "
               "cannot hold this batch; split it or use lsg_incident_features_dev", n_incidents, LSG_STATIC_MAX_INCIDENTS);
    return LSG_ERR_RANGE;
  }
  if (n_incidents == 0) return 0;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0) {
    static_err(errbuf, errbuf_len, "no current CUDA device (cudaGetDevice: %s)", cudaGetErrorString(e));
    return LSG_ERR_NODEVICE;
  }
  if (dev >= kStaticMaxDevices) {
    static_err(errbuf, errbuf_len, "device ordinal %d is beyond the %d devices this entry point keeps an address for",
               dev, kStaticMaxDevices);
    return LSG_ERR_RANGE;
  }
  unsigned long long *keys = nullptr;
  int sms = 0;
  {
    // set-up, once per device, with that device current: resolve the array's address on this GPU
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_keys_dev[dev]) {
      unsigned long long *p = nullptr;
      e = cudaGetSymbolAddress((void **)&p, g_static_incident_keys);
      if (e != cudaSuccess || !p) {
        static_err(errbuf, errbuf_len, "cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        return LSG_ERR_CUDA;
      }
      int n_sm = 0;
      e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
      if (e != cudaSuccess || n_sm <= 0) {
        static_err(errbuf, errbuf_len, "cudaDeviceGetAttribute failed on device %d: %s", dev, cudaGetErrorString(e));
        return LSG_ERR_CUDA;
      }
      g_static_sm_count[dev] = n_sm;
      g_static_keys_dev[dev] = p;
    }
    keys = g_static_keys_dev[dev];
    sms = g_static_sm_count[dev];
  }
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n = n_incidents * LSG_NSIG;
  // the array still holds the previous call's maxima: everything this call will read starts from "absent"
  e = cudaMemsetAsync(keys, 0, n * sizeof(unsigned long long), st);
  if (e != cudaSuccess) {
    static_err(errbuf, errbuf_len, "cudaMemsetAsync failed: %s", cudaGetErrorString(e));
    return LSG_ERR_CUDA;
  }
  auto blocks_for = [&](size_t items) {
    size_t b = (items + 255) / 256, cap = (size_t)sms * 8;
    return (unsigned)(b > cap ? cap : (b ? b : 1));
  };
  if (n_spans)
    k_incident_accumulate<<<blocks_for(n_spans * LSG_NSIG), 256, 0, st>>>(span_attr, span_incident, nullptr, n_spans,
                                                                        n_incidents, keys, nullptr);
  k_incident_finish<<<blocks_for(n), 256, 0, st>>>(keys, n, signals_out);
  e = cudaGetLastError();   // also reports (and clears) a sticky error left by an earlier, unrelated launch
  if (e != cudaSuccess) {
    static_err(errbuf, errbuf_len, "kernel launch failed (or an earlier launch on this device left an error): %s",
               cudaGetErrorString(e));
    return LSG_ERR_CUDA;
  }
  return 0;
}
