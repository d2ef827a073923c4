// stream.cu — the rows of SURVEY.md §8(f) that sit either side of the hot path:
//   F2  ring-buffer record decode   pkg/collector/ringbuf.go:42-52 (bpfEvent), :152-238 (decode, toProbeEvent,
//       signalFromType, convertValue) -> lsg_event records ready for the sort
//   F4  signal status               pkg/signals/generator.go:203-242 (signalStatus, threshold)
//   F1  retry-storm window counts   pkg/correlation/retry_storm.go:46-62,98-110 (Record + pruneOld) over a batch
#include "common.cuh"

namespace {

inline unsigned grid_for(lsg_ctx *ctx, size_t n, int block, int per_sm) {
  size_t blocks = (n + block - 1) / block;
  size_t cap = (size_t)ctx->sm_count * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  return (unsigned)blocks;
}

// signalStatus thresholds (warning, error) by signal index, generator.go:203-232
__constant__ double c_warn[LSG_NSIG] = {40, 2, 10, 80, 60, 2, 40, 5, 10, 50, 1, 1};
__constant__ double c_err[LSG_NSIG] = {120, 5, 25, 180, 160, 8, 120, 20, 50, 200, 3, 3};

__device__ __forceinline__ uint8_t status_of(unsigned sig, double v) {
  if (sig >= LSG_NSIG) return 0;          // default branch: "ok"
  if (v >= c_err[sig]) return 2;          // threshold(), generator.go:234-242
  if (v >= c_warn[sig]) return 1;
  return 0;
}

// llm_slo_event, ebpf/c/llm_slo_event.h:32-42: 40 bytes, packed, little endian
//   0 pid u32 | 4 tid u32 | 8 timestamp_ns u64 | 16 signal_type u32 | 20 value_ns u64 | 28 src_port u16 | 30 dst_port u16 |
//   32 dst_ip u32 | 36 errno i32
__global__ void __launch_bounds__(256)
k_decode_ringbuf(const uint32_t *__restrict__ raw, size_t n, uint32_t ident, uint32_t trace, int64_t ts_offset_ns,
                 lsg_event *__restrict__ out, uint8_t *__restrict__ status_out, int32_t *__restrict__ errno_out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const uint32_t *w = raw + i * 10;   // ten 32-bit words per record; records are only 4-byte aligned
    const uint32_t pid = w[0];
    const uint64_t ts = ((uint64_t)w[3] << 32) | w[2];
    const uint32_t st = w[4];
    const uint64_t value_ns = ((uint64_t)w[6] << 32) | w[5];
    const uint32_t ports = w[7], dst_ip = w[8];
    // signalFromType (ringbuf.go:199-226): kernel types 1..9 -> signal index in DefaultLikelihoods order
    unsigned sig = LSG_SIG_UNSUPPORTED;
    switch (st) {
      case 1: sig = 0; break; case 2: sig = 1; break; case 3: sig = 2; break; case 4: sig = 3; break; case 5: sig = 4; break;
      case 6: sig = 5; break; case 7: sig = 7; break; case 8: sig = 8; break; case 9: sig = 9; break;
      default: break;
    }
    // convertValue (ringbuf.go:228-238): counts and cpu_steal pass through, everything else ns -> ms
    const double v = (st == 2 || st == 6) ? (double)value_ns : (double)value_ns / 1e6;
    // ConnTuple is present when either port is non-zero (ringbuf.go:184-192); its identity is a 32-bit mix of
    // (src port, dst port, dst ip) — equal tuples get equal ids, distinct tuples collide with probability 2^-32
    uint32_t conn = 0;
    if (ports != 0) {
      uint32_t h = dst_ip * 0x9e3779b1u ^ ports * 0x85ebca6bu;
      h ^= h >> 15; h *= 0x2c1b3c6du; h ^= h >> 12;
      conn = h ? h : 1u;
    }
    U8 r;
    const uint64_t t = (uint64_t)((int64_t)ts + ts_offset_ns);
    const uint64_t vb = (uint64_t)__double_as_longlong(v);
    r.a = make_uint4((uint32_t)t, (uint32_t)(t >> 32), (uint32_t)vb, (uint32_t)(vb >> 32));
    r.b = make_uint4(ident, pid, conn, (trace << 4) | sig);
    st_v8(out + i, r);
    if (status_out) status_out[i] = status_of(sig, v);
    if (errno_out) errno_out[i] = (int32_t)w[9];
  }
}

__global__ void __launch_bounds__(256)
k_signal_status(const lsg_event *__restrict__ events, size_t n, uint8_t *__restrict__ status_out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const EventRec e = load_event(events + i);
    status_out[i] = status_of(e.trace_sig & 15u, e.value);
  }
}

// events of one pod are consecutive and time-ordered in `keys`; the count after Record(pod, ts) is the number of
// that pod's events with ts' in [ts - window, ts] that arrived no later (retry_storm.go:57-61, pruneOld :98-110)
__global__ void __launch_bounds__(256)
k_retry_counts(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ perm, size_t n, size_t n_live, uint32_t tbits,
               uint64_t window, uint32_t threshold, uint32_t *__restrict__ count_out, uint8_t *__restrict__ storm_out) {
  const uint64_t tmask = ((uint64_t)1 << tbits) - 1;
  for (size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (size_t)gridDim.x * blockDim.x) {
    uint32_t c = 0;
    if (j < n_live) {
      const uint64_t key = keys[j];
      const uint64_t t = key & tmask;
      const uint64_t target = (key & ~tmask) | (t > window ? t - window : 0);
      size_t lo = 0, hi = j;   // first position in [0, j] with key >= target
      while (lo < hi) {
        const size_t mid = lo + ((hi - lo) >> 1);
        if (keys[mid] < target) lo = mid + 1; else hi = mid;
      }
      c = (uint32_t)(j - lo + 1);
    }
    const uint32_t at = perm[j];
    count_out[at] = c;
    if (storm_out) storm_out[at] = c >= threshold && c != 0;
  }
}

template <typename T>
int stage_in(lsg_ctx *ctx, int slot, const T *host, size_t n, const T **dev) {
  *dev = nullptr;
  if (!host) return 0;
  void *p;
  LSG_TRY(lsg_ws(ctx, slot, n * sizeof(T), &p));
  if (n) LSG_CUDA(ctx, cudaMemcpyAsync(p, host, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  *dev = (const T *)p;
  return 0;
}

}  // namespace

extern "C" {

int lsg_decode_ringbuf_dev(lsg_ctx *ctx, const void *raw, size_t n, uint32_t ident, uint32_t trace, int64_t ts_offset_ns,
                           lsg_event *events_out, uint8_t *status_out, int32_t *errno_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!raw || !events_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  if (trace >= (1u << LSG_TRACE_BITS)) return lsg_fail(ctx, LSG_ERR_RANGE, "trace id %u does not fit %d bits", trace, LSG_TRACE_BITS);
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  if (n) {
    k_decode_ringbuf<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>((const uint32_t *)raw, n, ident, trace, ts_offset_ns,
                                                                        events_out, status_out, errno_out);
    LSG_LAUNCH_CHECK(ctx);
  }
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

int lsg_decode_ringbuf(lsg_ctx *ctx, const void *raw, size_t n, uint32_t ident, uint32_t trace, int64_t ts_offset_ns,
                       lsg_event *events_out, uint8_t *status_out, int32_t *errno_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!raw || !events_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  const uint8_t *d_raw;
  LSG_TRY(stage_in(ctx, WS_IN0, (const uint8_t *)raw, n * 40, &d_raw));
  void *d_ev, *d_st = nullptr, *d_er = nullptr;
  LSG_TRY(lsg_ws(ctx, WS_OUT0, n * sizeof(lsg_event), &d_ev));
  if (status_out) LSG_TRY(lsg_ws(ctx, WS_OUT1, n, &d_st));
  if (errno_out) LSG_TRY(lsg_ws(ctx, WS_OUT2, n * 4, &d_er));
  LSG_TRY(lsg_decode_ringbuf_dev(ctx, d_raw, n, ident, trace, ts_offset_ns, (lsg_event *)d_ev, (uint8_t *)d_st, (int32_t *)d_er));
  if (n) {
    LSG_CUDA(ctx, cudaMemcpyAsync(events_out, d_ev, n * sizeof(lsg_event), cudaMemcpyDeviceToHost, ctx->stream));
    if (status_out) LSG_CUDA(ctx, cudaMemcpyAsync(status_out, d_st, n, cudaMemcpyDeviceToHost, ctx->stream));
    if (errno_out) LSG_CUDA(ctx, cudaMemcpyAsync(errno_out, d_er, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  }
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

int lsg_signal_status_dev(lsg_ctx *ctx, const lsg_event *events, size_t n, uint8_t *status_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!events || !status_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  if (n) {
    k_signal_status<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(events, n, status_out);
    LSG_LAUNCH_CHECK(ctx);
  }
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

int lsg_signal_status(lsg_ctx *ctx, const lsg_event *events, size_t n, uint8_t *status_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!events || !status_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  const lsg_event *d_ev;
  LSG_TRY(stage_in(ctx, WS_EVENTS_IN, events, n, &d_ev));
  void *d_st;
  LSG_TRY(lsg_ws(ctx, WS_OUT0, n, &d_st));
  LSG_TRY(lsg_signal_status_dev(ctx, d_ev, n, (uint8_t *)d_st));
  if (n) LSG_CUDA(ctx, cudaMemcpyAsync(status_out, d_st, n, cudaMemcpyDeviceToHost, ctx->stream));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

int lsg_retry_storm_dev(lsg_ctx *ctx, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                        int64_t window_ns, int threshold, uint32_t *count_out, uint8_t *storm_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!events || !idents || !count_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  if (window_ns < 0) return lsg_fail(ctx, LSG_ERR_INVALID, "negative window");
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  const uint64_t *keys;
  const uint32_t *perm;
  KeySpec ks;
  size_t n_live;
  LSG_TRY(lsg_sort_signal_by_pod(ctx, events, n, idents, n_idents, 1u /* tcp_retransmits_total */, &keys, &perm, &ks, &n_live));
  if (n) {
    k_retry_counts<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(keys, perm, n, n_live, ks.tbits, (uint64_t)window_ns,
                                                                      (uint32_t)(threshold < 0 ? 0 : threshold), count_out, storm_out);
    LSG_LAUNCH_CHECK(ctx);
  }
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

int lsg_retry_storm(lsg_ctx *ctx, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                    int64_t window_ns, int threshold, uint32_t *count_out, uint8_t *storm_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if (n && (!events || !idents || !count_out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  const lsg_event *d_ev;
  const lsg_ident *d_id;
  LSG_TRY(stage_in(ctx, WS_EVENTS_IN, events, n, &d_ev));
  LSG_TRY(stage_in(ctx, WS_IDENTS, idents, n_idents, &d_id));
  void *d_cnt, *d_storm = nullptr;
  LSG_TRY(lsg_ws(ctx, WS_OUT0, n * 4, &d_cnt));
  if (storm_out) LSG_TRY(lsg_ws(ctx, WS_OUT1, n, &d_storm));
  LSG_TRY(lsg_retry_storm_dev(ctx, d_ev, n, d_id, n_idents, window_ns, threshold, (uint32_t *)d_cnt, (uint8_t *)d_storm));
  if (n) {
    LSG_CUDA(ctx, cudaMemcpyAsync(count_out, d_cnt, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (storm_out) LSG_CUDA(ctx, cudaMemcpyAsync(storm_out, d_storm, n, cudaMemcpyDeviceToHost, ctx->stream));
  }
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

}  // extern "C"
