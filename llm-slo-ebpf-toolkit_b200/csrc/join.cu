// join.cu — K-join: Correlator.EnrichAttributes for every span of a batch
// (pkg/otel/processor/ebpfcorrelator/correlator.go:50-125, processor.go:29-49) against events that are
// already in (join group, quantised time) order.
//
// The reference visits every (span, signal) pair.  Here a span visits only the events of its own join
// group whose quantised time lies inside the outer window — a contiguous range of the sorted array,
// found by a warp-cooperative 32-ary search — and applies the complete four-tier Match()
// (pkg/correlation/dns.go:50-76) with exact nanosecond arithmetic to each of them.  The reference's
// candidate order (confidence desc, |dt| asc, then input order: sort.SliceStable, correlator.go:92-99)
// is a total order on (tier, |dt|, input index); every lane keeps its best F under that order and the
// warp merges them, so neither the scan order nor the time quantisation can change the result.
//
// One warp per span; spans are processed in key order, so the warps of a CTA read overlapping event
// ranges (L1/L2 reuse).  Outputs are written at the span's input position.
#include <cmath>

#include "common.cuh"

namespace {

struct JoinParams {
  const lsg_span *spans;
  const uint32_t *span_perm;
  size_t n_spans;
  const lsg_event *ev;       // sorted records
  const uint32_t *ev_perm;   // sorted position -> input index
  const void *ev_keys;
  size_t n_live;
  KeySpec ks;
  int64_t ts_max_q;          // largest quantised time of the key space, 2^tbits - 1
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
};

// first position in [0, n) whose key is >= target (n if none); all lanes of the warp take part
template <typename KeyT>
__device__ __forceinline__ size_t warp_lower_bound(const KeyT *__restrict__ keys, size_t n, KeyT target, unsigned lane) {
  size_t lo = 0, hi = n;
  while (lo < hi) {
    const size_t len = hi - lo;
    const size_t step = (len + 32) / 33;
    const size_t probe = lo + (size_t)(lane + 1) * step - 1;
    const bool in = probe < hi;
    bool ge = true;
    if (in) ge = __ldg(keys + probe) >= target;
    const unsigned b = __ballot_sync(0xffffffffu, ge);
    if (b == 0) {
      lo = lo + 32 * step;
    } else {
      const unsigned f = __ffs(b) - 1;
      const size_t pf = lo + (size_t)(f + 1) * step - 1;
      hi = pf < hi ? pf : hi;
      lo = lo + (size_t)f * step;
      if (lo > hi) lo = hi;
    }
  }
  return lo;
}

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

template <typename KeyT, int F>
__global__ void __launch_bounds__(256)
k_join(const JoinParams p) {
  __shared__ unsigned long long s_cnt[5];
  if (threadIdx.x < 5) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31;
  const size_t warps_per_grid = (size_t)gridDim.x * (blockDim.x >> 5);
  const KeyT *keys = reinterpret_cast<const KeyT *>(p.ev_keys);
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  unsigned long long w_matched = 0, w_low = 0, w_drop = 0, w_cand = 0, w_kept = 0;

  for (size_t q = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); q < p.n_spans; q += warps_per_grid) {
    const size_t si = p.span_perm ? p.span_perm[q] : q;
    const SpanRec s = load_span(p.spans + si);
    TopF<F> top;
    top.init();
    unsigned n_cand = 0, n_match = 0, n_low = 0;
    bool ok_ident = s.ident < p.n_idents;
    if (!ok_ident && lane == 0) atomicOr((unsigned int *)&p.counters[5], 1u);
    lsg_ident sid;
    sid.pod = sid.node = sid.service = sid.grp = 0;
    if (ok_ident) sid = load_ident(p.idents, s.ident);
    if (sid.grp != 0 && s.ts != LSG_TS_ZERO && p.n_live != 0) {
      // quantised window [qlo, qhi] in the key space of the events, clamped to the times that exist
      const int64_t W = p.window_ns;
      const int64_t t_lo = s.ts < INT64_MIN + W ? INT64_MIN : s.ts - W;
      const int64_t t_hi = s.ts > INT64_MAX - W ? INT64_MAX : s.ts + W;
      const int64_t ts_min = p.ks.ts_min;
      // offsets from ts_min as exact unsigned magnitudes; a window that ends before the first event or
      // starts after the last representable quantum is empty
      const uint64_t qlo = t_lo <= ts_min ? 0 : (((uint64_t)t_lo - (uint64_t)ts_min) >> p.ks.shift);
      if (t_hi >= ts_min && qlo <= (uint64_t)p.ts_max_q) {
        uint64_t qhi = ((uint64_t)t_hi - (uint64_t)ts_min) >> p.ks.shift;
        if (qhi > (uint64_t)p.ts_max_q) qhi = (uint64_t)p.ts_max_q;
        const KeyT klo = (KeyT)(((uint64_t)sid.grp << p.ks.tbits) | qlo);
        const KeyT khi = (KeyT)(((uint64_t)sid.grp << p.ks.tbits) | qhi);
        const size_t a = warp_lower_bound<KeyT>(keys, p.n_live, klo, lane);
        const size_t b = warp_lower_bound<KeyT>(keys, p.n_live, (KeyT)(khi + 1), lane);
        for (size_t j = a + lane; j < b; j += 32) {
          const EventRec e = load_event(p.ev + j);
          const uint64_t adt = lsg_absdiff(s.ts, e.ts);
          if (adt > (uint64_t)W) continue;   // outer window, dns.go:54-56 (quantisation only widens the scan)
          const lsg_ident eid = load_ident(p.idents, e.ident);
          const int tier = lsg_match_tier(adt, s.trace, s.pid, s.conn, sid, e.trace_sig >> 4, e.pid, e.conn, eid);
          if (tier == 0) continue;           // Unmatched (counted analytically by the caller)
          n_match++;
          if (lsg_tier_conf(tier) < p.threshold) { n_low++; continue; }   // correlator.go:83-86
          n_cand++;
          top.insert(((uint64_t)tier << 56) | adt, p.ev_perm[j], (uint32_t)j);
        }
      }
    }
    // warp totals
    for (int o = 16; o; o >>= 1) {
      n_cand += __shfl_xor_sync(0xffffffffu, n_cand, o);
      n_match += __shfl_xor_sync(0xffffffffu, n_match, o);
      n_low += __shfl_xor_sync(0xffffffffu, n_low, o);
    }
    // merge the lanes' best-F lists into the span's best `fanout` (reference order)
    uint32_t my_idx = LSG_NONE, my_pos = 0xffffffffu;
    unsigned my_tier = 0;
    unsigned first_tier = 0;
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
    // outputs: kept candidates in order (correlator.go:101-104)
    if ((int)lane < p.fanout) {
      p.cand_event[si * (size_t)p.fanout + lane] = my_idx;
      p.cand_tier[si * (size_t)p.fanout + lane] = (uint8_t)my_tier;
    }
    // per-attribute max over the kept candidates (correlator.go:106-112): lane s owns signal s
    double val = nan;
    if (lane < LSG_NSIG && p.attr_base) val = p.attr_base[si * LSG_NSIG + lane];
    for (int r = 0; r < kept; ++r) {
      const uint32_t ps = __shfl_sync(0xffffffffu, my_pos, r);
      const lsg_event *e = p.ev + ps;
      const double v = e->value;                      // same address in every lane: one broadcast load
      const unsigned sig = e->trace_sig & 15u;
      if (sig == lane && (isnan(val) || v > val)) val = v;
    }
    if (lane < LSG_NSIG) p.attr_out[si * LSG_NSIG + lane] = val;
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
  if (lane == 0) {
    if (w_matched) atomicAdd(&s_cnt[0], w_matched);
    if (w_low) atomicAdd(&s_cnt[1], w_low);
    if (w_drop) atomicAdd(&s_cnt[2], w_drop);
    if (w_cand) atomicAdd(&s_cnt[3], w_cand);
    if (w_kept) atomicAdd(&s_cnt[4], w_kept);
  }
  __syncthreads();
  if (threadIdx.x < 5 && s_cnt[threadIdx.x]) atomicAdd(&p.counters[threadIdx.x], s_cnt[threadIdx.x]);
}

}  // namespace

int lsg_launch_join(lsg_ctx *ctx, const JoinArgs &a) {
  LSG_CUDA(ctx, cudaMemsetAsync(a.counters, 0, 8 * sizeof(uint64_t), ctx->stream));
  if (a.n_spans == 0) return 0;
  JoinParams p;
  p.spans = a.spans; p.span_perm = a.span_perm; p.n_spans = a.n_spans;
  p.ev = a.ev.rec; p.ev_perm = a.ev.perm; p.ev_keys = a.ev.keys; p.n_live = a.ev.n_live;
  p.ks = a.ev.key;
  p.ts_max_q = (int64_t)((((uint64_t)1 << a.ev.key.tbits) - 1));
  p.idents = a.idents; p.n_idents = (uint32_t)a.n_idents;
  p.window_ns = a.window_ns; p.threshold = a.threshold; p.fanout = a.fanout;
  p.attr_base = a.attr_base; p.cand_event = a.cand_event; p.cand_tier = a.cand_tier; p.attr_out = a.attr_out;
  p.conf_out = a.conf_out; p.retrieval_out = a.retrieval_out; p.n_cand_out = a.n_cand_out;
  p.counters = (unsigned long long *)a.counters;
  size_t blocks = (a.n_spans + 7) / 8;
  size_t cap = (size_t)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  const bool k32 = a.ev.key.bits <= 31;
  ProfScope ps(ctx, PK_JOIN);
  if (a.fanout <= 3) {
    if (k32) k_join<uint32_t, 3><<<(unsigned)blocks, 256, 0, ctx->stream>>>(p);
    else k_join<uint64_t, 3><<<(unsigned)blocks, 256, 0, ctx->stream>>>(p);
  } else {
    if (k32) k_join<uint32_t, LSG_MAX_FANOUT><<<(unsigned)blocks, 256, 0, ctx->stream>>>(p);
    else k_join<uint64_t, LSG_MAX_FANOUT><<<(unsigned)blocks, 256, 0, ctx->stream>>>(p);
  }
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}
