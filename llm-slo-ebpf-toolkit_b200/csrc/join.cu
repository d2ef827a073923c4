// join.cu — K-join: Correlator.EnrichAttributes for every span of a batch
// (pkg/otel/processor/ebpfcorrelator/correlator.go:50-125, processor.go:29-49) against events that are
// already in (join group, quantised time) order.
//
// The reference visits every (span, signal) pair.  Here a span visits only the events of its own join
// group whose quantised time lies inside the outer window — a contiguous range [a, b) of the sorted
// columns — and applies the complete four-tier Match() (pkg/correlation/dns.go:50-76) with exact
// nanosecond arithmetic to each of them.  The reference's candidate order (confidence desc, |dt| asc,
// then input order: sort.SliceStable, correlator.go:92-99) is a total order on (tier, |dt|, input
// index); every lane keeps its best F under that order and the warp merges them, so neither the scan
// order nor the time quantisation can change the result.
//
// Two kernels:
//   k_span_ranges  one thread per span (in key order): binary search of its [a, b) in the sorted keys.
//   k_join_tiles   a CTA takes 128 consecutive spans.  Their ranges overlap heavily (each event is inside
//                  ~35 spans' windows at C2), so the CTA stages the union of the ranges — five columns,
//                  24 B/event — into shared memory once with TMA bulk copies (cp.async.bulk + mbarrier)
//                  and its 8 warps scan their spans' windows out of shared memory, one warp per span.
//                  HBM/L2 traffic drops from ~11 KB to ~0.5 KB per span.  A span whose own range does
//                  not fit a tile is scanned straight from global memory by the same code.
// Outputs are written at the span's input position; optionally the per-incident maxima (A7 glue) are
// accumulated from the epilogue with one atomicMax per present attribute.
#include <cmath>

#include "common.cuh"

namespace {

constexpr int JT_THREADS = 256;
constexpr int JT_WARPS = JT_THREADS / 32;
constexpr int JT_CH = 128;     // spans per chunk
constexpr int JT_CAP = 2048;   // events per shared-memory tile

struct JoinParams {
  const lsg_span *spans;
  const uint32_t *span_perm;
  size_t n_spans;
  EventColumns ev;
  const uint32_t *ev_perm;   // sorted position -> input index
  const uint2 *ranges;       // per span in key order: [a, b) in sorted positions
  const lsg_ident *idents;
  uint32_t n_idents;
  int64_t window_ns;
  double threshold;
  int fanout;
  const double *attr_base;
  uint32_t *cand_event;
  uint8_t *cand_tier;
  double *attr_out;
  double *conf_out;
  double *retrieval_out;
  uint32_t *n_cand_out;
  unsigned long long *counters;  // [0] matched [1] low_confidence [2] fanout_dropped [3] candidates [4] kept [5] bad ident
  unsigned long long *inc_keys;
  size_t n_incidents;
};

// ---- per-span event range --------------------------------------------------------------------------
template <typename KeyT>
__device__ __forceinline__ uint32_t lower_bound_keys(const KeyT *__restrict__ keys, uint32_t n, KeyT target) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    if (__ldg(keys + mid) < target) lo = mid + 1; else hi = mid;
  }
  return lo;
}

template <typename KeyT>
__global__ void __launch_bounds__(256)
k_span_ranges(const lsg_span *__restrict__ spans, const uint32_t *__restrict__ span_perm, size_t n_spans,
              const KeyT *__restrict__ keys, uint32_t n_live, KeySpec ks, const lsg_ident *__restrict__ idents,
              uint32_t n_idents, int64_t W, uint2 *__restrict__ ranges, unsigned long long *__restrict__ counters) {
  const uint64_t ts_max_q = ((uint64_t)1 << ks.tbits) - 1;
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_spans; q += (size_t)gridDim.x * blockDim.x) {
    const size_t si = span_perm ? span_perm[q] : q;
    const SpanRec s = load_span(spans + si);
    uint2 r = make_uint2(n_live, n_live);
    if (s.ident >= n_idents) {
      atomicOr((unsigned int *)&counters[5], 1u);
    } else {
      const uint32_t grp = load_ident(idents, s.ident).grp;
      if (grp != 0 && s.ts != LSG_TS_ZERO && n_live != 0) {
        // The range is a function of the span's SORT KEY, not of its exact time: spans are processed in key
        // order, and the tiles below rely on a and b being non-decreasing in that order.  So take the span's
        // quantum [t0, t1] (its time clamped into the key space exactly as k_span_keys does) and search for
        // [t0 - W, t1 + W]: a superset of the exact window (the scan re-checks |dt| <= W in nanoseconds),
        // identical for spans that share a key and monotone in the key.
        const int64_t key_ts_max = (int64_t)((uint64_t)ks.ts_min + (ts_max_q << ks.shift));
        const int64_t tc = s.ts < ks.ts_min ? ks.ts_min : (s.ts > key_ts_max ? key_ts_max : s.ts);
        const uint64_t qs = ((uint64_t)tc - (uint64_t)ks.ts_min) >> ks.shift;
        const int64_t t0 = (int64_t)((uint64_t)ks.ts_min + (qs << ks.shift));
        const int64_t t1 = (int64_t)((uint64_t)t0 + (((uint64_t)1 << ks.shift) - 1));
        const int64_t t_lo = t0 < INT64_MIN + W ? INT64_MIN : t0 - W;
        const int64_t t_hi = t1 > INT64_MAX - W ? INT64_MAX : t1 + W;
        const uint64_t qlo = t_lo <= ks.ts_min ? 0 : (((uint64_t)t_lo - (uint64_t)ks.ts_min) >> ks.shift);
        uint64_t qhi = ((uint64_t)t_hi - (uint64_t)ks.ts_min) >> ks.shift;   // t_hi >= t1 >= ts_min
        if (qhi > ts_max_q) qhi = ts_max_q;
        const KeyT klo = (KeyT)(((uint64_t)grp << ks.tbits) | qlo);
        const KeyT khi = (KeyT)(((uint64_t)grp << ks.tbits) | qhi);
        r.x = lower_bound_keys<KeyT>(keys, n_live, klo);
        r.y = lower_bound_keys<KeyT>(keys, n_live, (KeyT)(khi + 1));
      }
    }
    ranges[q] = r;
  }
}

// ---- best-F candidates under the reference's order ----------------------------------------------------
template <int F>
struct TopF {
  uint64_t k1[F];   // (tier << 56) | |dt|
  uint32_t idx[F];  // input index of the event
  uint32_t pos[F];  // sorted position of the event
  __device__ __forceinline__ void init() {
#pragma unroll
    for (int i = 0; i < F; ++i) { k1[i] = ~0ull; idx[i] = 0xffffffffu; pos[i] = 0xffffffffu; }
  }
  __device__ __forceinline__ void insert(uint64_t k, uint32_t ix, uint32_t ps) {
#pragma unroll
    for (int i = 0; i < F; ++i) {
      const bool less = k < k1[i] || (k == k1[i] && ix < idx[i]);
      if (less) {
        uint64_t tk = k1[i]; k1[i] = k; k = tk;
        uint32_t ti = idx[i]; idx[i] = ix; ix = ti;
        uint32_t tp = pos[i]; pos[i] = ps; ps = tp;
      }
    }
  }
  __device__ __forceinline__ void pop() {
#pragma unroll
    for (int i = 0; i + 1 < F; ++i) { k1[i] = k1[i + 1]; idx[i] = idx[i + 1]; pos[i] = pos[i + 1]; }
    k1[F - 1] = ~0ull; idx[F - 1] = 0xffffffffu; pos[F - 1] = 0xffffffffu;
  }
};

// ---- TMA bulk copy + mbarrier (global -> shared, completion counted in bytes) ---------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t phase) {
  uint32_t ok;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(phase) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct JoinSmem {
  int64_t ts[JT_CAP];
  uint32_t ident[JT_CAP];
  uint32_t pid[JT_CAP];
  uint32_t conn[JT_CAP];
  uint32_t trsig[JT_CAP];
  uint4 span[JT_CH][2];
  uint32_t a[JT_CH], b[JT_CH], sidx[JT_CH];
  unsigned long long cnt[5];
  uint64_t bar;
};

// the five columns a scan reads, either the shared-memory tile (base = first staged position) or global memory (base 0)
struct Cols {
  const int64_t *ts;
  const uint32_t *ident, *pid, *conn, *trsig;
  uint32_t base;
};

// One span, one warp: scan [a, b), select, write.  Counters go to the caller's per-lane accumulators.
template <int F>
__device__ __forceinline__ void process_span(const JoinParams &p, const Cols c, const SpanRec s, const size_t si, const uint32_t a,
                                             const uint32_t b, const unsigned lane, unsigned long long &w_matched,
                                             unsigned long long &w_low, unsigned long long &w_drop, unsigned long long &w_cand,
                                             unsigned long long &w_kept) {
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  TopF<F> top;
  top.init();
  unsigned n_cand = 0, n_match = 0, n_low = 0;
  if (a < b) {
    const lsg_ident sid = load_ident(p.idents, s.ident);
    const uint64_t W = (uint64_t)p.window_ns;
    for (uint32_t j = a + lane; j < b; j += 32) {
      const uint32_t o = j - c.base;
      const uint64_t adt = lsg_absdiff(s.ts, c.ts[o]);
      if (adt > W) continue;   // outer window, dns.go:54-56 (the quantised range is only a superset)
      const uint32_t trsig = c.trsig[o];
      int tier = 0;
      if (s.trace != 0 && s.trace == (trsig >> 4)) {
        tier = LSG_TIER_TRACE;                                       // dns.go:58-60
      } else if (adt <= (uint64_t)(500 * kMs)) {                     // tiers 2-4 all need |dt| <= 500 ms
        const uint32_t e_ident = c.ident[o];
        lsg_ident eid = sid;
        if (e_ident != s.ident) eid = load_ident(p.idents, e_ident);
        tier = lsg_match_tier(adt, 0, s.pid, s.conn, sid, 0, c.pid[o], c.conn[o], eid);
      }
      if (tier == 0) continue;   // Unmatched (counted analytically by the caller)
      n_match++;
      if (lsg_tier_conf(tier) < p.threshold) { n_low++; continue; }   // correlator.go:83-86
      n_cand++;
      top.insert(((uint64_t)tier << 56) | adt, p.ev_perm[j], j);
    }
  }
  for (int o = 16; o; o >>= 1) {
    n_cand += __shfl_xor_sync(0xffffffffu, n_cand, o);
    n_match += __shfl_xor_sync(0xffffffffu, n_match, o);
    n_low += __shfl_xor_sync(0xffffffffu, n_low, o);
  }
  // merge the lanes' best-F lists into the span's best `fanout` (reference order)
  uint32_t my_idx = LSG_NONE, my_pos = 0xffffffffu;
  unsigned my_tier = 0, first_tier = 0;
  const int want = p.fanout < F ? p.fanout : F;
  int kept = 0;
  if (n_cand != 0) {
    for (int r = 0; r < want; ++r) {
      uint64_t mk = top.k1[0];
      uint32_t mi = top.idx[0];
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        const uint64_t ok = __shfl_xor_sync(0xffffffffu, mk, o);
        const uint32_t oi = __shfl_xor_sync(0xffffffffu, mi, o);
        if (ok < mk || (ok == mk && oi < mi)) { mk = ok; mi = oi; }
      }
      if (mk == ~0ull) break;
      const bool mine = top.k1[0] == mk && top.idx[0] == mi;
      const unsigned owner = __ffs(__ballot_sync(0xffffffffu, mine)) - 1;
      const uint32_t ps = __shfl_sync(0xffffffffu, top.pos[0], owner);
      if (mine) top.pop();
      if ((int)lane == r) { my_idx = mi; my_pos = ps; my_tier = (unsigned)(mk >> 56); }
      if (r == 0) first_tier = (unsigned)(mk >> 56);
      kept++;
    }
  }
  // kept candidates in order (correlator.go:101-104)
  if ((int)lane < p.fanout) {
    p.cand_event[si * (size_t)p.fanout + lane] = my_idx;
    p.cand_tier[si * (size_t)p.fanout + lane] = (uint8_t)my_tier;
  }
  // per-attribute max over the kept candidates (correlator.go:106-112): lane s owns signal s
  double val = nan;
  if (lane < LSG_NSIG && p.attr_base) val = p.attr_base[si * LSG_NSIG + lane];
  for (int r = 0; r < kept; ++r) {
    const uint32_t ps = __shfl_sync(0xffffffffu, my_pos, r);
    const double v = p.ev.val[ps];                     // same address in every lane: one broadcast load
    const unsigned sig = p.ev.trsig[ps] & 15u;
    if (sig == lane && (isnan(val) || v > val)) val = v;
  }
  if (lane < LSG_NSIG) {
    p.attr_out[si * LSG_NSIG + lane] = val;
    if (p.inc_keys && s.incident != LSG_NONE && s.incident < p.n_incidents && !isnan(val))
      atomicMax(&p.inc_keys[(size_t)s.incident * LSG_NSIG + lane], f64_key(val));   // A7 glue: max per incident and signal
  }
  // llm.ebpf.correlation_confidence = max confidence = the first kept candidate's (correlator.go:113-118)
  const double v3 = __shfl_sync(0xffffffffu, val, 3), v4 = __shfl_sync(0xffffffffu, val, 4);
  if (lane == 0) {
    p.conf_out[si] = kept ? lsg_tier_conf((int)first_tier) : 0.0;
    if (p.retrieval_out) {   // DecomposeRetrieval, correlator.go:179-194: dns + connect + tls in that order
      double total = 0.0;
      if (!isnan(val)) total += val;
      if (!isnan(v3)) total += v3;
      if (!isnan(v4)) total += v4;
      p.retrieval_out[si] = total;
    }
    if (p.n_cand_out) p.n_cand_out[si] = n_cand;
    w_matched += n_match;
    w_low += n_low;
    w_cand += n_cand;
    w_kept += kept;
    if (n_cand > (unsigned)p.fanout) w_drop += n_cand - p.fanout;   // correlator.go:101-103
  }
}

__device__ __forceinline__ SpanRec span_from_halves(const uint4 sa, const uint4 sb) {
  SpanRec s;
  s.ts = (int64_t)(((uint64_t)sa.y << 32) | sa.x);
  s.ident = sa.z; s.pid = sa.w; s.conn = sb.x; s.trace = sb.y; s.incident = sb.z;
  return s;
}

template <int F>
__global__ void __launch_bounds__(JT_THREADS)
k_join_tiles(const __grid_constant__ JoinParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  JoinSmem &sm = *reinterpret_cast<JoinSmem *>(smem_raw);
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 5) sm.cnt[tid] = 0;
  if (tid == 0) mbar_init(&sm.bar, 1);
  __syncthreads();
  uint32_t phase = 0;
  unsigned long long w_matched = 0, w_low = 0, w_drop = 0, w_cand = 0, w_kept = 0;
  const size_t n_chunks = (p.n_spans + JT_CH - 1) / JT_CH;

  for (size_t chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
    const size_t q0 = chunk * JT_CH;
    const int nq = (int)min((size_t)JT_CH, p.n_spans - q0);
    // stage the chunk's spans: input index, range, record (two 16-byte halves per span)
    if ((int)tid < nq) {
      const uint2 r = p.ranges[q0 + tid];
      sm.a[tid] = r.x;
      sm.b[tid] = r.y;
      sm.sidx[tid] = p.span_perm ? p.span_perm[q0 + tid] : (uint32_t)(q0 + tid);
    }
    {
      const int k = tid >> 1, half = tid & 1;
      if (k < nq) {
        const size_t si = p.span_perm ? p.span_perm[q0 + k] : q0 + k;
        sm.span[k][half] = ld_nc_v4(reinterpret_cast<const uint4 *>(p.spans + si) + half);
      }
    }
    __syncthreads();
    int i0 = 0;
    while (i0 < nq) {
      const uint32_t A = sm.a[i0] & ~3u;   // 16-byte aligned start for every column
      // spans of the chunk whose range still ends inside a tile that starts at A (b is non-decreasing)
      const bool fits = (int)tid >= i0 && (int)tid < nq && sm.b[tid] - A <= (uint32_t)(JT_CAP - 4);
      const int n_fit = __syncthreads_count(fits);
      if (n_fit == 0) {
        // this span's own range is larger than a tile: scan it from global memory
        if (warp == 0) {
          const SpanRec s = span_from_halves(sm.span[i0][0], sm.span[i0][1]);
          const Cols g{p.ev.ts, p.ev.ident, p.ev.pid, p.ev.conn, p.ev.trsig, 0u};
          process_span<F>(p, g, s, sm.sidx[i0], sm.a[i0], sm.b[i0], lane, w_matched, w_low, w_drop, w_cand, w_kept);
        }
        i0 += 1;
        continue;
      }
      const int i1 = i0 + n_fit;
      const uint32_t Bend = (sm.b[i1 - 1] + 3u) & ~3u;
      const uint32_t count = Bend > A ? Bend - A : 0;   // <= JT_CAP, multiple of 4; the columns are padded past n
      if (count) {
        if (tid == 0) {
          mbar_expect_tx(&sm.bar, count * 24u);
          bulk_g2s(sm.ts, p.ev.ts + A, count * 8u, &sm.bar);
          bulk_g2s(sm.trsig, p.ev.trsig + A, count * 4u, &sm.bar);
          bulk_g2s(sm.ident, p.ev.ident + A, count * 4u, &sm.bar);
          bulk_g2s(sm.pid, p.ev.pid + A, count * 4u, &sm.bar);
          bulk_g2s(sm.conn, p.ev.conn + A, count * 4u, &sm.bar);
        }
        mbar_wait(&sm.bar, phase);
        phase ^= 1;
      }
      const Cols c{sm.ts, sm.ident, sm.pid, sm.conn, sm.trsig, A};
      for (int k = i0 + (int)warp; k < i1; k += JT_WARPS) {
        const SpanRec s = span_from_halves(sm.span[k][0], sm.span[k][1]);
        process_span<F>(p, c, s, sm.sidx[k], sm.a[k], sm.b[k], lane, w_matched, w_low, w_drop, w_cand, w_kept);
      }
      __syncthreads();   // every warp is done with the tile before the next bulk copy overwrites it
      i0 = i1;
    }
    __syncthreads();     // and with the chunk's span staging before the next chunk
  }
  if (lane == 0) {
    if (w_matched) atomicAdd(&sm.cnt[0], w_matched);
    if (w_low) atomicAdd(&sm.cnt[1], w_low);
    if (w_drop) atomicAdd(&sm.cnt[2], w_drop);
    if (w_cand) atomicAdd(&sm.cnt[3], w_cand);
    if (w_kept) atomicAdd(&sm.cnt[4], w_kept);
  }
  __syncthreads();
  if (tid < 5 && sm.cnt[tid]) atomicAdd(&p.counters[tid], sm.cnt[tid]);
}

}  // namespace

int lsg_launch_join(lsg_ctx *ctx, const JoinArgs &a) {
  LSG_CUDA(ctx, cudaMemsetAsync(a.counters, 0, 8 * sizeof(uint64_t), ctx->stream));
  if (a.n_spans == 0) return 0;
  void *rv = nullptr;
  LSG_TRY(lsg_ws(ctx, WS_RANGES, a.n_spans * sizeof(uint2), &rv));
  ProfScope ps(ctx, PK_JOIN);
  const uint32_t n_live = (uint32_t)a.ev.n_live;
  {
    size_t blocks = (a.n_spans + 255) / 256, cap = (size_t)ctx->sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (a.ev.key.bits <= 31)
      k_span_ranges<uint32_t><<<(unsigned)blocks, 256, 0, ctx->stream>>>(a.spans, a.span_perm, a.n_spans, (const uint32_t *)a.ev.keys,
                                                                      n_live, a.ev.key, a.idents, (uint32_t)a.n_idents,
                                                                      a.window_ns, (uint2 *)rv, (unsigned long long *)a.counters);
    else
      k_span_ranges<uint64_t><<<(unsigned)blocks, 256, 0, ctx->stream>>>(a.spans, a.span_perm, a.n_spans, (const uint64_t *)a.ev.keys,
                                                                      n_live, a.ev.key, a.idents, (uint32_t)a.n_idents,
                                                                      a.window_ns, (uint2 *)rv, (unsigned long long *)a.counters);
    LSG_LAUNCH_CHECK(ctx);
  }
  JoinParams p;
  p.spans = a.spans; p.span_perm = a.span_perm; p.n_spans = a.n_spans;
  p.ev = a.ev.col; p.ev_perm = a.ev.perm; p.ranges = (const uint2 *)rv;
  p.idents = a.idents; p.n_idents = (uint32_t)a.n_idents;
  p.window_ns = a.window_ns; p.threshold = a.threshold; p.fanout = a.fanout;
  p.attr_base = a.attr_base; p.cand_event = a.cand_event; p.cand_tier = a.cand_tier; p.attr_out = a.attr_out;
  p.conf_out = a.conf_out; p.retrieval_out = a.retrieval_out; p.n_cand_out = a.n_cand_out;
  p.counters = (unsigned long long *)a.counters;
  p.inc_keys = a.inc_keys; p.n_incidents = a.n_incidents;
  const size_t n_chunks = (a.n_spans + JT_CH - 1) / JT_CH;
  size_t blocks = n_chunks, cap = (size_t)ctx->sm_count * 4;
  if (blocks > cap) blocks = cap;
  const size_t smem = sizeof(JoinSmem);
  if (a.fanout <= 3) {
    LSG_CUDA(ctx, cudaFuncSetAttribute(k_join_tiles<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_join_tiles<3><<<(unsigned)blocks, JT_THREADS, smem, ctx->stream>>>(p);
  } else {
    LSG_CUDA(ctx, cudaFuncSetAttribute(k_join_tiles<LSG_MAX_FANOUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_join_tiles<LSG_MAX_FANOUT><<<(unsigned)blocks, JT_THREADS, smem, ctx->stream>>>(p);
  }
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}
