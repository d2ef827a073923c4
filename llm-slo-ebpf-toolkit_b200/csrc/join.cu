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
// Kernels:
//   k_span_ranges  one thread per span (in key order): its outer range [a, b) (|dt| <= W, the only place a
//                  trace-id match can be), its inner range [a2, b2) (|dt| <= min(W, 500 ms), the only place
//                  tiers 2-4 can be), the whole time quanta inside the exact inner window and the +-250 ms part
//                  of the inner range: reads of the counting sort's bin table, or binary searches in the
//                  sorted keys on the radix path.
//   k_ident_class_*  per identity row (pod, class of its (service, node) pair), so that a (span, event) pair
//                  compares two words instead of loading two identity rows.
//   k_join_tiles   a CTA takes 128 consecutive spans.  Their ranges overlap heavily (each event is inside
//                  ~35 spans' windows at C2), so the CTA stages the union of the ranges into shared memory
//                  once, as six columns (time, pod, class, pid, conn, trace|signal: 28 B/event).  FUSED (what the
//                  pipeline runs): there is no sorted copy of the events; each thread reads its permutation entries
//                  and then the records themselves (256-bit loads, four in flight), writes the columns and probes
//                  the trace-id hash while the record is in registers.  Unfused (LSG_JOIN_SORTED_COPY): TMA bulk
//                  copies (cp.async.bulk + mbarrier) of the sorted columns, then the identity column is rewritten
//                  to (pod, class) in place and a second pass probes the hash.  Tier 1 is that tile-local hash join
//                  (every tile event probes a hash of the chunk's span trace ids once).  Tiers 2-4 are decided
//                  by two threads per span out of shared memory; where every event of the tile carries one
//                  class and the service+node tier is below the threshold, that tier is counted (whole quanta
//                  + exact edge tests - trace-id matches - tier 2/3 matches) and only the +-250 ms part is
//                  enumerated.  Matches are noted in two bytes and ordered afterwards in registers.
//   k_join_big     the spans the tile kernel leaves out (range larger than a tile, overflowing note list, time
//                  outside the key space, fanout > 8): one warp per span straight from global memory.
// Outputs are written at the span's input position; optionally the per-incident maxima (A7 glue) are
// accumulated from the epilogue with one atomicMax per present attribute.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <type_traits>

#include "common.cuh"

namespace {

// Shape of the tile kernel.  The macros exist for scratch/build_variant.sh (the same library with other constants, to
// measure them); what was measured on C2 / C3, join ms (profiles/r02_summary.md):
//   256 threads, 1792 events, 48 notes, 3 CTAs/SM   0.600 / 72.3     <- (with two records in flight per loader thread: 0.587)
//   256 threads, 1920 events, 32 notes, 3 CTAs/SM   0.600 / 68.0     (0.592: more spans overflow their note list and are
//   256 threads, 2048 events, 24 notes, 3 CTAs/SM   0.604 / 67.3      redone by k_join_big, 23 us per step instead of 5)
//   256 threads, 2816 events, 2 CTAs/SM             0.700 / 74.7
//   256 threads, 1152 events, 4 CTAs/SM (64 regs)   0.699
//   128 threads, 1152 events, 5 CTAs/SM             0.668
//   128 threads,  896 events, 6 CTAs/SM             0.851
// More resident CTAs do not pay for a smaller tile: every tile gathers the +-window halo of its spans again.
#ifndef LSG_JT_THREADS
#define LSG_JT_THREADS 256
#endif
#ifndef LSG_TM_CAP
#define LSG_TM_CAP 1792
#endif
#ifndef LSG_JT_OCC
#define LSG_JT_OCC 3
#endif
#ifndef LSG_JT_DEPTH
#define LSG_JT_DEPTH 2   // records a thread of the loader has in flight (measured on C2 / C3: 1: 0.608 / 67.0, 2: 0.592 / 66.8, 4: 0.605 / 68.6, 8: 0.625 / 70.1)
#endif
constexpr int JT_THREADS = LSG_JT_THREADS;
constexpr int JT_WARPS = JT_THREADS / 32;
constexpr int JT_CH = JT_THREADS / 2;     // spans per chunk: two threads per span

struct JoinParams {
  const lsg_span *spans;
  const uint32_t *span_perm;
  size_t n_spans;
  EventColumns ev;           // events in key order as columns; all NULL when no sorted copy was made (FUSED kernels)
  const lsg_event *events_in;   // the input records (what ev_perm indexes)
  const uint32_t *ev_perm;   // sorted position -> input index
  const uint4 *ranges;       // per span in key order: outer [x, y) and inner [z, w) in sorted positions
  const uint4 *ranges2;      // whole-quanta core [x, y) of the inner window and the +-250 ms part [z, w) of the inner range
  const lsg_ident *idents;
  const uint2 *ident_pc;     // per identity row: (pod, service+node class)
  uint32_t n_idents;
  int64_t window_ns;
  double threshold;
  int fanout;
  const double *attr_base;
  uint32_t *cand_event;
  uint8_t *cand_tier;
  double *attr_out;          // may be NULL (not wanted)
  double *conf_out;          // may be NULL
  double *cand_value;        // optional compact form: value and signal of the kept candidates, [n_spans * fanout]
  uint8_t *cand_sig;
  double *retrieval_out;
  uint32_t *n_cand_out;
  unsigned long long *counters;  // [0] matched [1] low_confidence [2] fanout_dropped [3] candidates [4] kept [5] bad ident
  unsigned long long *inc_keys;
  size_t n_incidents;
  lsg_feature *feat;             // optional: count / sum / log2 sketch per (incident, signal) next to the maxima
  uint32_t seg_max;              // k_join_ring: spans per segment at most
};

// one span's enriched value of one signal into the incident's features (the same bins as k_incident_accumulate)
__device__ __forceinline__ void feat_add(lsg_feature *feat, size_t slot, double v) {
  atomicAdd(&feat[slot].count, 1u);
  atomicAdd(&feat[slot].sum, v);
  int bin = 0;
  if (v > 0) {
    int ex;
    frexp(v, &ex);
    bin = ex - 1 + 4;
    bin = bin < 0 ? 0 : (bin > LSG_SKETCH_BINS - 1 ? LSG_SKETCH_BINS - 1 : bin);
  }
  atomicAdd(&feat[slot].sketch[bin], 1u);
}

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
              const KeyT *__restrict__ keys, const uint32_t *__restrict__ bins, uint32_t n_live, KeySpec ks,
              const lsg_ident *__restrict__ idents,
              uint32_t n_idents, int64_t W, uint4 *__restrict__ ranges, uint4 *__restrict__ ranges2,
              unsigned long long *__restrict__ counters) {
  const uint64_t ts_max_q = ((uint64_t)1 << ks.tbits) - 1;
  // counting sort: only the placed events have a permutation entry, and bins[dead] is their number.  A span that can
  // match nothing gets the empty range at that position, so a tile that holds it never reads past the written part
  // of the permutation (the host's n_live is only an upper bound on that path).
  if (bins) n_live = bins[(uint32_t)ks.dead];
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_spans; q += (size_t)gridDim.x * blockDim.x) {
    const size_t si = span_perm ? span_perm[q] : q;
    const SpanRec s = load_span(spans + si);
    uint4 r = make_uint4(n_live, n_live, n_live, n_live);
    uint4 r2 = r;   // x, y: the events of the quanta that lie wholly inside the exact inner window; z, w: superset of +-250 ms
    if (s.ident >= n_idents) {
      atomicOr((unsigned int *)&counters[5], 1u);
    } else {
      const uint32_t grp = load_ident(idents, s.ident).grp;
      // (a group beyond the key space can only come from a key hint that does not hold: the batch is void, stay in bounds)
      if (grp != 0 && ((uint64_t)grp << ks.tbits) < ks.dead && s.ts != LSG_TS_ZERO && n_live != 0) {
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
        auto search = [&](const int64_t w, uint32_t lo_pos, uint32_t hi_pos, uint32_t *pa, uint32_t *pb) {
          const int64_t t_lo = t0 < INT64_MIN + w ? INT64_MIN : t0 - w;
          const int64_t t_hi = t1 > INT64_MAX - w ? INT64_MAX : t1 + w;
          const uint64_t qlo = t_lo <= ks.ts_min ? 0 : (((uint64_t)t_lo - (uint64_t)ks.ts_min) >> ks.shift);
          uint64_t qhi = ((uint64_t)t_hi - (uint64_t)ks.ts_min) >> ks.shift;   // t_hi >= t1 >= ts_min
          if (qhi > ts_max_q) qhi = ts_max_q;
          const KeyT klo = (KeyT)(((uint64_t)grp << ks.tbits) | qlo);
          const KeyT khi = (KeyT)(((uint64_t)grp << ks.tbits) | qhi);
          if (bins) {   // counting sort: bins[k] is the start of key k's run
            *pa = bins[(uint32_t)klo];
            *pb = bins[(uint32_t)khi + 1];
          } else {
            *pa = lo_pos + lower_bound_keys<KeyT>(keys + lo_pos, hi_pos - lo_pos, klo);
            *pb = lo_pos + lower_bound_keys<KeyT>(keys + lo_pos, hi_pos - lo_pos, (KeyT)(khi + 1));
          }
        };
        search(W, 0, n_live, &r.x, &r.y);
        const int64_t Wi = W < 500 * kMs ? W : 500 * kMs;          // tiers 2-4: dns.go:62-73
        search(Wi, r.x, r.y, &r.z, &r.w);                           // inside the outer range: a short search
        // a span whose own time lies outside the events' time range was clamped into the key space: its inner
        // range is still a superset, but the tile kernel's 32-bit time deltas are not safe for it -> k_join_big
        if (tc != s.ts) {
          r.z = r.w = 0xffffffffu;
        } else {
          // positions of whole quanta, from the span's exact time this time (these need not be monotone): events of the
          // quanta in [qa, qb] are inside |dt| <= Wi without looking at them; tiers 2-3 need |dt| <= 250 ms (dns.go:61-68)
          auto pos_of_quantum = [&](uint64_t qv) -> uint32_t {   // first event of the group at or after quantum qv
            if (qv > ts_max_q) return r.w;
            const KeyT k = (KeyT)(((uint64_t)grp << ks.tbits) | qv);
            if (bins) return bins[(uint32_t)k];
            return r.z + lower_bound_keys<KeyT>(keys + r.z, r.w - r.z, k);
          };
          const uint64_t quantum = (uint64_t)1 << ks.shift;
          const uint64_t off = (uint64_t)s.ts - (uint64_t)ks.ts_min;
          const uint64_t qa = off > (uint64_t)Wi ? (off - (uint64_t)Wi + quantum - 1) >> ks.shift : 0;   // first quantum starting at or after ts - Wi
          const uint64_t qb1 = (off + (uint64_t)Wi + 1) >> ks.shift;                                         // first quantum NOT ending at or before ts + Wi
          if (qa < qb1) {
            r2.x = max(r.z, min(r.w, pos_of_quantum(qa)));
            r2.y = max(r2.x, min(r.w, pos_of_quantum(qb1)));
          } else {
            r2.x = r2.y = r.w;
          }
          const uint64_t W2 = (uint64_t)(Wi < 250 * kMs ? Wi : 250 * kMs);
          const uint64_t qc = off > W2 ? (off - W2) >> ks.shift : 0;
          const uint64_t qd1 = ((off + W2) >> ks.shift) + 1;
          r2.z = max(r.z, min(r.w, pos_of_quantum(qc)));
          r2.w = max(r2.z, min(r.w, pos_of_quantum(qd1)));
        }
      }
    }
    ranges[q] = r;
    ranges2[q] = r2;
  }
}

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

// generic-proxy accesses of a shared-memory buffer (the previous tile's reads) ordered before an async-proxy write to it
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr int JT_LIST = 64;    // per-warp candidate list (entries); compacted to the best `fanout` when it fills

struct Cand {
  uint64_t k1;    // (tier << 56) | |dt|: the reference's order, correlator.go:92-99
  uint32_t idx;   // input index of the event: ties keep input order (sort.SliceStable)
  uint32_t pos;   // sorted position of the event
};

// What the warp-per-span scan reads: the events in key order, either as the sorted columns or, when no sorted copy
// was made (the tile kernel gathers its own tiles), as the input records behind the permutation.
struct ColsSorted {
  const int64_t *ts_;
  const double *val_;
  const uint32_t *ident_, *pid_, *conn_, *trsig_, *perm_;
  __device__ __forceinline__ int64_t ts(uint32_t o) const { return ts_[o]; }
  __device__ __forceinline__ double val(uint32_t o) const { return val_[o]; }
  __device__ __forceinline__ uint32_t ident(uint32_t o) const { return ident_[o]; }
  __device__ __forceinline__ uint32_t pid(uint32_t o) const { return pid_[o]; }
  __device__ __forceinline__ uint32_t conn(uint32_t o) const { return conn_[o]; }
  __device__ __forceinline__ uint32_t trsig(uint32_t o) const { return trsig_[o]; }
  __device__ __forceinline__ uint32_t perm(uint32_t o) const { return perm_[o]; }
  // four consecutive trace|signal words, o a multiple of 4 (the columns are padded past the last event)
  __device__ __forceinline__ uint4 trsig4(uint32_t o, uint32_t) const { return *reinterpret_cast<const uint4 *>(trsig_ + o); }
};
struct ColsIndirect {
  const lsg_event *ev_;
  const uint32_t *perm_;
  __device__ __forceinline__ const lsg_event *at(uint32_t o) const { return ev_ + perm_[o]; }
  __device__ __forceinline__ int64_t ts(uint32_t o) const { return at(o)->ts_ns; }
  __device__ __forceinline__ double val(uint32_t o) const { return at(o)->value; }
  __device__ __forceinline__ uint32_t ident(uint32_t o) const { return at(o)->ident; }
  __device__ __forceinline__ uint32_t pid(uint32_t o) const { return at(o)->pid; }
  __device__ __forceinline__ uint32_t conn(uint32_t o) const { return at(o)->conn; }
  __device__ __forceinline__ uint32_t trsig(uint32_t o) const { return at(o)->trace_sig; }
  __device__ __forceinline__ uint32_t perm(uint32_t o) const { return perm_[o]; }
  __device__ __forceinline__ uint4 trsig4(uint32_t o, uint32_t end) const {   // positions at or past `end` are not looked at
    uint4 q = make_uint4(0, 0, 0, 0);
    if (o < end) q.x = trsig(o);
    if (o + 1 < end) q.y = trsig(o + 1);
    if (o + 2 < end) q.z = trsig(o + 2);
    if (o + 3 < end) q.w = trsig(o + 3);
    return q;
  }
};

__device__ __forceinline__ bool cand_less(uint64_t ka, uint32_t ia, uint64_t kb, uint32_t ib) {
  return ka < kb || (ka == kb && ia < ib);
}

// Best `want` entries of list[0, n) in order; lane r ends up holding the r-th.  Returns how many there are.
__device__ __forceinline__ int select_best(const Cand *list, const int n, const int want, const unsigned lane, uint64_t &my_k,
                                           uint32_t &my_idx, uint32_t &my_pos) {
  // every lane owns entries lane and lane + 32 (n <= JT_LIST = 64)
  Cand e0{~0ull, 0xffffffffu, 0xffffffffu}, e1 = e0;
  if ((int)lane < n) e0 = list[lane];
  if ((int)lane + 32 < n) e1 = list[lane + 32];
  my_k = ~0ull; my_idx = LSG_NONE; my_pos = 0xffffffffu;
  int kept = 0;
  for (int r = 0; r < want; ++r) {
    const bool second = cand_less(e1.k1, e1.idx, e0.k1, e0.idx);
    const uint64_t k = second ? e1.k1 : e0.k1;
    const uint32_t ix = second ? e1.idx : e0.idx, ps = second ? e1.pos : e0.pos;
    // lexicographic warp minimum of (k.hi, k.lo, idx): three REDUX steps
    const uint32_t hi = (uint32_t)(k >> 32), lo = (uint32_t)k;
    const uint32_t mhi = __reduce_min_sync(0xffffffffu, hi);
    const uint32_t mlo = __reduce_min_sync(0xffffffffu, hi == mhi ? lo : 0xffffffffu);
    const uint32_t mi = __reduce_min_sync(0xffffffffu, (hi == mhi && lo == mlo) ? ix : 0xffffffffu);
    const uint64_t mk = ((uint64_t)mhi << 32) | mlo;
    if (mk == ~0ull) break;
    const bool mine = k == mk && ix == mi;
    const unsigned owner = __ffs(__ballot_sync(0xffffffffu, mine)) - 1;
    const uint32_t wpos = __shfl_sync(0xffffffffu, ps, owner);
    if (mine) {
      if (second) e1.k1 = ~0ull, e1.idx = 0xffffffffu; else e0.k1 = ~0ull, e0.idx = 0xffffffffu;
    }
    if ((int)lane == r) { my_k = mk; my_idx = mi; my_pos = wpos; }
    kept++;
  }
  return kept;
}

// One span, one warp: scan, select, write.  Counters go to the caller's per-lane accumulators.
template <typename Cols>
__device__ __forceinline__ void process_span(const JoinParams &p, const Cols c, Cand *list, const SpanRec s, const size_t si,
                                             const uint4 rg_in, const unsigned lane, unsigned long long &w_matched,
                                             unsigned long long &w_low, unsigned long long &w_drop, unsigned long long &w_cand,
                                             unsigned long long &w_kept) {
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  const unsigned lt_mask = (1u << lane) - 1u;
  uint4 rg = rg_in;
  if (rg.z == 0xffffffffu) { rg.z = rg.x; rg.w = rg.y; }   // no inner range recorded: the outer one is a superset
  const uint64_t W = (uint64_t)p.window_ns;
  unsigned n_cand = 0, n_match = 0, n_low = 0;
  int n_list = 0;   // warp-uniform
  uint64_t my_k;
  uint32_t my_idx, my_pos;

  // append the lanes that have a candidate to the warp's list (warp-uniform control flow around it)
  auto append = [&](const bool has, const uint64_t k1, const uint32_t o, const uint32_t j) {
    const unsigned m = __ballot_sync(0xffffffffu, has);
    if (m == 0) return;
    if (n_list + 32 > JT_LIST) {   // keep only what can still win
      __syncwarp();
      const int kept = select_best(list, n_list, p.fanout, lane, my_k, my_idx, my_pos);
      __syncwarp();
      if ((int)lane < kept) list[lane] = Cand{my_k, my_idx, my_pos};
      n_list = kept;
      __syncwarp();
    }
    if (has) list[n_list + __popc(m & lt_mask)] = Cand{k1, c.perm(o), j};
    n_list += __popc(m);
    n_cand += __popc(m);   // warp-uniform running total
  };

  if (rg.x < rg.y) {
    // ---- tier 1 (dns.go:58-60): a trace-id match can sit anywhere in the outer window.  Only the trace|signal
    //      column is read, four events per lane and load; the time is looked at for hits only.
    if (s.trace != 0) {
      const uint32_t v0 = rg.x & ~3u, v_end = rg.y, v_beg = rg.x;
      for (uint32_t vb = v0; vb < v_end; vb += 128) {
        const uint32_t v = vb + lane * 4;
        uint4 q = make_uint4(0, 0, 0, 0);
        if (v < v_end) q = c.trsig4(v, v_end);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const uint32_t o = v + e;
          bool hit = o >= v_beg && o < v_end && (w[e] >> 4) == s.trace;
          uint64_t adt = 0;
          if (hit) {
            adt = lsg_absdiff(s.ts, c.ts(o));
            hit = adt <= W;                                   // outer window, dns.go:54-56
          }
          if (__any_sync(0xffffffffu, hit)) {
            n_match += __popc(__ballot_sync(0xffffffffu, hit));
            const bool low = hit && 1.0 < p.threshold;       // confidence 1.0 below the threshold (threshold > 1)
            n_low += __popc(__ballot_sync(0xffffffffu, low));
            append(hit && !low, ((uint64_t)LSG_TIER_TRACE << 56) | adt, o, o);
          }
        }
      }
    }
    // ---- tiers 2-4 (dns.go:61-73) need |dt| <= 500 ms: the inner range only; trace-id matches were taken above
    if (rg.z < rg.w) {
      const lsg_ident sid = load_ident(p.idents, s.ident);
      const uint64_t Wi = W < (uint64_t)(500 * kMs) ? W : (uint64_t)(500 * kMs);
      for (uint32_t jb = rg.z; jb < rg.w; jb += 32) {
        const uint32_t j = jb + lane;
        int tier = 0;
        uint64_t adt = 0;
        uint32_t o = 0;
        if (j < rg.w) {
          o = j;
          adt = lsg_absdiff(s.ts, c.ts(o));
          const uint32_t trsig = c.trsig(o);
          if (adt <= Wi && !(s.trace != 0 && s.trace == (trsig >> 4))) {
            const uint32_t e_ident = c.ident(o);
            lsg_ident eid = sid;
            if (e_ident != s.ident) eid = load_ident(p.idents, e_ident);
            tier = lsg_match_tier(adt, 0, s.pid, s.conn, sid, 0, c.pid(o), c.conn(o), eid);
          }
        }
        const bool is_m = tier != 0;
        const unsigned mm = __ballot_sync(0xffffffffu, is_m);
        if (mm) {
          const bool low = is_m && lsg_tier_conf(tier) < p.threshold;   // correlator.go:83-86
          n_match += __popc(mm);
          n_low += __popc(__ballot_sync(0xffffffffu, low));
          append(is_m && !low, ((uint64_t)tier << 56) | adt, o, j);
        }
      }
    }
  }
  // ---- the span's best `fanout` candidates in the reference's order (correlator.go:92-104)
  __syncwarp();
  const int kept = select_best(list, n_list, p.fanout, lane, my_k, my_idx, my_pos);
  __syncwarp();
  const unsigned my_tier = my_idx == LSG_NONE ? 0u : (unsigned)(my_k >> 56);
  const unsigned first_tier = __shfl_sync(0xffffffffu, my_tier, 0);
  if ((int)lane < p.fanout) {
    p.cand_event[si * (size_t)p.fanout + lane] = my_idx;
    p.cand_tier[si * (size_t)p.fanout + lane] = (uint8_t)my_tier;
  }
  // per-attribute max over the kept candidates (correlator.go:106-112): lane s owns signal s.
  // lane r < kept fetches its own candidate's value and signal (independent loads, one latency), then broadcasts
  double val = nan;
  if (lane < LSG_NSIG && p.attr_base) val = p.attr_base[si * LSG_NSIG + lane];
  double my_val = 0.0;
  unsigned my_sig = 0xffu;
  if ((int)lane < kept) {
    my_val = c.val(my_pos);
    my_sig = c.trsig(my_pos) & 15u;
  }
  if ((int)lane < p.fanout) {
    if (p.cand_value) p.cand_value[si * (size_t)p.fanout + lane] = my_val;
    if (p.cand_sig) p.cand_sig[si * (size_t)p.fanout + lane] = (uint8_t)my_sig;
  }
  for (int r = 0; r < kept; ++r) {
    const double v = __shfl_sync(0xffffffffu, my_val, r);
    const unsigned sig = __shfl_sync(0xffffffffu, my_sig, r);
    if (sig == lane && (isnan(val) || v > val)) val = v;
  }
  if (lane < LSG_NSIG) {
    if (p.attr_out) p.attr_out[si * LSG_NSIG + lane] = val;
    if (p.inc_keys && s.incident != LSG_NONE && s.incident < p.n_incidents && !isnan(val)) {
      atomicMax(&p.inc_keys[(size_t)s.incident * LSG_NSIG + lane], f64_key(val));   // A7 glue: max per incident and signal
      if (p.feat) feat_add(p.feat, (size_t)s.incident * LSG_NSIG + lane, val);
    }
  }
  // llm.ebpf.correlation_confidence = max confidence = the first kept candidate's (correlator.go:113-118)
  const double v3 = __shfl_sync(0xffffffffu, val, 3), v4 = __shfl_sync(0xffffffffu, val, 4);
  if (lane == 0) {
    if (p.conf_out) p.conf_out[si] = kept ? lsg_tier_conf((int)first_tier) : 0.0;
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

// ---- the tile kernel -------------------------------------------------------------------------------------
// A warp that serves one span pays a vote or ballot for every rare event, and scanning a span's whole outer
// window for trace-id matches repeats, for every span, work that is an equality join.  So per tile:
//   tier 1 (dns.go:58-60)  the chunk's span trace ids sit in a 256-slot shared-memory hash; every event of the
//                          tile is looked at ONCE by one thread and probes it.  A same-trace pair is a match iff
//                          |dt| <= W, and such an event is inside the span's range, hence inside the tile.
//   tiers 2-4 (dns.go:61-73) need |dt| <= 500 ms: two threads per span (lanes 2k, 2k+1) each walk half of the
//                          span's inner range with plain per-thread code; time deltas fit 32 bits there.
// A match is only noted — two bytes (tier, tile offset) in the span's list — and the lists are ordered
// afterwards in registers, when every lane is doing the same thing.  A span whose list overflows (or whose range
// does not fit a tile) is left to k_join_big.
constexpr int TM_CAP = LSG_TM_CAP;    // events per tile (28 B each; three CTAs per SM)
#ifndef LSG_TM_LIST
#define LSG_TM_LIST 48
#endif
constexpr int TM_LIST = LSG_TM_LIST;     // noted candidates per span
constexpr int TM_MAX_F = 8;      // largest fanout the tile kernel keeps in registers
constexpr int TM_HASH = JT_THREADS;    // hash slots for the chunk's span trace ids (<= 128 keys)
constexpr uint16_t TM_END = 0xffffu;

// ---- tiers 2-4 for one thread's share of a span's inner range (both forms of the tile kernel) -------------------
// The events are the tile's shared-memory columns; `slot(g)` turns the loop variable into a column index (identity for
// the fixed tile, `g mod capacity` for the ring).  The decision per event is predicate arithmetic on five loaded words.
// A match at or above the threshold is not noted inside the loop (a divergent branch with a shared-memory atomic that
// some lane of the warp takes in nearly every iteration): the loop runs in blocks of 32 events and collects the matches
// of a block as bit masks, which are noted afterwards, every lane at its own pace.  With T32 the time deltas and the
// tier limits are 32-bit (|dt| < 2^31 ns inside an inner range).
struct PairSpan {
  int64_t ts;
  uint32_t s_lo, pod, cls, pid, conn, no_trace;
  bool pod_ok, pid_ok, conn_ok, full;
};
struct TileCols {
  const int64_t *ts;
  const uint32_t *pod, *cls, *pid, *conn, *trsig;
};

template <bool T32, bool WITH4, int LIST, typename Slot>
__device__ __forceinline__ void pair_scan(const TileCols c, const uint32_t beg, const uint32_t end, const Slot slot, const PairSpan &sp,
                                          const uint64_t Wi, const int cut, uint32_t *n_note_k, uint16_t *note_k /* stride JT_CH */,
                                          unsigned &n_match, unsigned &n_low, unsigned &n_unlisted) {
  using dt_t = typename std::conditional<T32, uint32_t, uint64_t>::type;
  const dt_t wi = (dt_t)Wi, w100 = (dt_t)(100 * kMs), w250 = (dt_t)(250 * kMs);
  const uint32_t *ts_lo = reinterpret_cast<const uint32_t *>(c.ts);   // little endian: low words at even indices
  for (uint32_t b = beg; b < end; b += 32) {
    const uint32_t be = min(b + 32u, end);
    uint32_t keep = 0, is2 = 0, is3 = 0;   // bit j: event b + j is a candidate to note / of tier 2 / of tier 3
    uint32_t bit = 1u;
    for (uint32_t g = b; g < be; ++g, bit <<= 1) {
      const uint32_t sl = slot(g);
      const uint32_t e_trsig = c.trsig[sl], e_pod = c.pod[sl], e_pid = c.pid[sl], e_conn = c.conn[sl];
      dt_t ad;
      if (T32) {
        const int32_t d = (int32_t)(ts_lo[2 * sl] - sp.s_lo);
        ad = (dt_t)(uint32_t)(d < 0 ? -d : d);
      } else {
        ad = (dt_t)lsg_absdiff(sp.ts, c.ts[sl]);
      }
      const bool in = ad <= wi, in100 = ad <= w100, in250 = ad <= w250;
      const bool same_pod = sp.pod_ok & (e_pod == sp.pod);
      const bool t2 = same_pod & sp.pid_ok & (e_pid == sp.pid) & in100;        // dns.go:61-64
      const bool t3 = same_pod & sp.conn_ok & (e_conn == sp.conn) & in250;     // dns.go:65-68
      bool t4 = false;
      if (WITH4) t4 = c.cls[sl] == sp.cls;                                      // dns.go:69-73
      const int tier = t2 ? LSG_TIER_POD_PID : t3 ? LSG_TIER_POD_CONN : LSG_TIER_SVC_NODE;
      // inside min(W, 500 ms), not a trace-id match (those belong to the hash pass), and some tier applies
      const bool m = in & ((e_trsig >> 4) != sp.no_trace) & (t2 | t3 | t4);
      n_match += m;
      const bool low = m & (tier > cut);                                       // correlator.go:83-86
      n_low += low;
      const bool k = m & !low;
      keep |= k ? bit : 0u;
      is2 |= t2 ? bit : 0u;
      is3 |= t3 ? bit : 0u;
    }
    if (sp.full) {
      n_unlisted += __popc(keep);
    } else {
      while (keep) {
        const int j = __ffs(keep) - 1;
        keep &= keep - 1;
        const int tier = (is2 >> j) & 1u ? LSG_TIER_POD_PID : (is3 >> j) & 1u ? LSG_TIER_POD_CONN : LSG_TIER_SVC_NODE;
        const uint32_t at = atomicAdd(n_note_k, 1u);
        if (at < (uint32_t)LIST) note_k[at * JT_CH] = (uint16_t)((tier << 12) | slot(b + j));
      }
    }
  }
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
};

// ---- what the tile kernels do with a span once its matches are noted ---------------------------------------------
// Called by every thread of the warp (shuffles inside); lanes 2k and 2k + 1 serve one span, h = lane & 1.
//   order  lane h takes the span's notes of parity h, then the odd lane hands its best F to the even lane: the
//          reference's order (tier, |dt|, input index), correlator.go:92-99
//   write  the even lane owns the span's candidates, confidence and counters (correlator.go:101-118); the twelve
//          attribute columns are split between the two lanes (six each), which both read the kept candidates' values
//          and signals; per-incident maxima (and optional features) by atomics: the A7 glue
// `pos_of` turns a note's column index into the event's sorted position, `slot_of` a sorted position into its column
// index.  A span whose list overflowed (or whose time was clamped into the key space) is listed for k_join_big.
struct TileTotals {
  unsigned long long matched = 0, low = 0, drop = 0, cand = 0, kept = 0;
};

template <int F, bool FUSED, int LIST, typename PosOf, typename SlotOf>
__device__ __forceinline__ void span_epilogue(const JoinParams &p, const TileCols c, const uint16_t *note_k /* stride JT_CH */,
                                              const uint32_t n_noted, const uint32_t n_tr, const bool live, const bool clamped,
                                              const SpanRec &s, const size_t si, const uint32_t span_pos, const unsigned h,
                                              const unsigned lane, const int cut, unsigned n_match, unsigned n_low, unsigned n_unlisted,
                                              const PosOf pos_of, const SlotOf slot_of, uint32_t *big_list, unsigned int *big_count,
                                              TileTotals &w) {
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  TopF<F> top;
  top.init();
  uint32_t n_cand = 0;
  bool redo = false;
  if (live) {
    n_cand = n_noted;
    redo = n_cand > (uint32_t)LIST || clamped;
    if (!redo) {
#pragma unroll 2
      for (uint32_t i = h; i < n_cand; i += 2) {
        const uint32_t e = note_k[i * JT_CH];
        const uint32_t sl = e & 0xfffu, tier = e >> 12;
        const uint32_t g = pos_of(sl);
        const uint64_t adt = lsg_absdiff(s.ts, c.ts[sl]);
        top.insert(((uint64_t)tier << 56) | adt, p.ev_perm[g], g);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < F; ++r) {
    const uint64_t ok = __shfl_down_sync(0xffffffffu, top.k1[r], 1);
    const uint32_t oi = __shfl_down_sync(0xffffffffu, top.idx[r], 1);
    const uint32_t op = __shfl_down_sync(0xffffffffu, top.pos[r], 1);
    if (!h && oi != 0xffffffffu) top.insert(ok, oi, op);
  }
  n_match += __shfl_down_sync(0xffffffffu, n_match, 1);
  n_low += __shfl_down_sync(0xffffffffu, n_low, 1);
  n_unlisted += __shfl_down_sync(0xffffffffu, n_unlisted, 1);
  n_cand += n_unlisted;   // candidates that were counted without being noted
  if (!live) return;
  if (redo) {
    if (!h) big_list[atomicAdd(big_count, 1u)] = span_pos;   // more candidates than the list holds
    return;
  }
  const unsigned pair = 3u << (lane & 30u);
  double vals[F];
  unsigned sigs[F];
  unsigned present = 0;
  int kept = 0;
#pragma unroll
  for (int r = 0; r < F; ++r) {
    const bool mine = r < p.fanout && top.idx[r] != 0xffffffffu;
    const uint32_t pos = __shfl_sync(pair, mine ? top.pos[r] : 0xffffffffu, lane & 30u);   // the even lane's final list
    const bool has = pos != 0xffffffffu;
    // the candidate's value: from the sorted column, or from its input record when there is no sorted copy
    const uint32_t in_idx = __shfl_sync(pair, top.idx[r], lane & 30u);
    vals[r] = !has ? 0.0 : FUSED ? p.events_in[in_idx].value : p.ev.val[pos];
    sigs[r] = has ? (c.trsig[slot_of(pos)] & 15u) : 0xffu;
    kept += has;
    present |= sigs[r] < LSG_NSIG ? 1u << sigs[r] : 0u;
    if (!h && r < p.fanout) {
      p.cand_event[si * (size_t)p.fanout + r] = has ? top.idx[r] : LSG_NONE;      // correlator.go:101-104
      p.cand_tier[si * (size_t)p.fanout + r] = has ? (uint8_t)(top.k1[r] >> 56) : (uint8_t)0;
      if (p.cand_value) p.cand_value[si * (size_t)p.fanout + r] = vals[r];
      if (p.cand_sig) p.cand_sig[si * (size_t)p.fanout + r] = (uint8_t)sigs[r];
    }
  }
  // per-attribute max over the kept candidates (correlator.go:106-112), one attribute at a time from registers
  double v0 = nan, v3 = nan, v4 = nan;
  const bool to_incident = p.inc_keys && s.incident != LSG_NONE && s.incident < p.n_incidents;
#pragma unroll
  for (int cc = 0; cc < LSG_NSIG / 2; ++cc) {
    const unsigned col = h * (LSG_NSIG / 2) + cc;
    double v = p.attr_base ? p.attr_base[si * LSG_NSIG + col] : nan;
    if ((present >> col) & 1u) {
#pragma unroll
      for (int r = 0; r < F; ++r)
        if (sigs[r] == col && (isnan(v) || vals[r] > v)) v = vals[r];
    }
    if (p.attr_out) p.attr_out[si * LSG_NSIG + col] = v;
    if (to_incident && !isnan(v)) {
      atomicMax(&p.inc_keys[(size_t)s.incident * LSG_NSIG + col], f64_key(v));   // A7 glue: max per incident and signal
      if (p.feat) feat_add(p.feat, (size_t)s.incident * LSG_NSIG + col, v);
    }
    if (cc == 0) v0 = v;   // on the even lane: dns (0), connect (3), tls (4)
    if (cc == 3) v3 = v;
    if (cc == 4) v4 = v;
  }
  if (!h) {
    n_match += n_tr;
    if (cut < 1) n_low += n_tr;
    // llm.ebpf.correlation_confidence = the first kept candidate's confidence (correlator.go:113-118)
    if (p.conf_out) p.conf_out[si] = kept ? lsg_tier_conf((int)(top.k1[0] >> 56)) : 0.0;
    if (p.retrieval_out) {   // DecomposeRetrieval, correlator.go:179-194: dns + connect + tls in that order
      double total = 0.0;
      if (!isnan(v0)) total += v0;
      if (!isnan(v3)) total += v3;
      if (!isnan(v4)) total += v4;
      p.retrieval_out[si] = total;
    }
    if (p.n_cand_out) p.n_cand_out[si] = n_cand;
    w.matched += n_match;
    w.low += n_low;
    w.cand += n_cand;
    w.kept += kept;
    if (n_cand > (uint32_t)p.fanout) w.drop += n_cand - p.fanout;   // correlator.go:101-103
  }
}

// block-level sum of the per-thread totals into the five join counters: lanes -> warp -> block -> one atomic each
__device__ __forceinline__ void flush_totals(TileTotals w, unsigned long long *block_cnt /* shared, zeroed */, unsigned long long *counters) {
  for (int o = 16; o; o >>= 1) {
    w.matched += __shfl_xor_sync(0xffffffffu, w.matched, o);
    w.low += __shfl_xor_sync(0xffffffffu, w.low, o);
    w.drop += __shfl_xor_sync(0xffffffffu, w.drop, o);
    w.cand += __shfl_xor_sync(0xffffffffu, w.cand, o);
    w.kept += __shfl_xor_sync(0xffffffffu, w.kept, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (w.matched) atomicAdd(&block_cnt[0], w.matched);
    if (w.low) atomicAdd(&block_cnt[1], w.low);
    if (w.drop) atomicAdd(&block_cnt[2], w.drop);
    if (w.cand) atomicAdd(&block_cnt[3], w.cand);
    if (w.kept) atomicAdd(&block_cnt[4], w.kept);
  }
  __syncthreads();
  if (threadIdx.x < 5 && block_cnt[threadIdx.x]) atomicAdd(&counters[threadIdx.x], block_cnt[threadIdx.x]);
}

struct TileSmem {
  int64_t ts[TM_CAP];
  uint32_t pod[TM_CAP];             // staged as the identity row, rewritten in place to the row's pod id
  uint32_t cls[TM_CAP];             // the row's (service, node) class
  uint32_t pid[TM_CAP];
  uint32_t conn[TM_CAP];
  uint32_t trsig[TM_CAP];
  uint4 span[JT_CH][2];
  uint4 rngbuf[2][JT_CH];           // this chunk's ranges and, arriving by bulk copy meanwhile, the next chunk's
  uint4 rng2[JT_CH];
  uint32_t sidx[JT_CH];
  uint16_t note[TM_LIST][JT_CH];    // entry-major: consecutive spans, consecutive halfwords
  uint32_t n_note[JT_CH];           // candidates noted per span (may exceed TM_LIST: then the span is redone by k_join_big)
  uint32_t n_tr[JT_CH];             // trace-id matches per span (tier 1), candidates or not
  uint32_t n_tr_in[JT_CH];          // those of them inside the inner window
  uint32_t cls_min, cls_max;        // over the tile's events: equal when the tile holds one (service, node) class
  uint32_t hkey[TM_HASH];           // trace id, 0 = empty
  uint16_t hhead[TM_HASH];          // first span of the chain for that trace id
  uint16_t hnext[JT_CH];            // next span with the same trace id
  unsigned long long cnt[5];
  uint64_t bar;
  uint64_t bar_rng;
};

__device__ __forceinline__ uint32_t trace_slot(uint32_t tr) { return (tr * 2654435761u) >> (TM_HASH == 256 ? 24 : TM_HASH == 128 ? 25 : 23); }

// FUSED: there is no sorted copy of the events; the CTA gathers its tile's records through the permutation itself
// (random 32-byte reads, issued four deep per thread), writing the columns and the (pod, class) words in one go.
template <int F, bool T32, bool FUSED>
__global__ void __launch_bounds__(JT_THREADS, LSG_JT_OCC)
k_join_tiles(const __grid_constant__ JoinParams p, uint32_t *__restrict__ big_list, unsigned int *__restrict__ big_count) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TileSmem &sm = *reinterpret_cast<TileSmem *>(smem_raw);
  const unsigned tid = threadIdx.x, lane = tid & 31;
  if (tid < 5) sm.cnt[tid] = 0;
  if (tid == 0) { mbar_init(&sm.bar, 1); mbar_init(&sm.bar_rng, 1); }
  __syncthreads();
  uint32_t phase = 0, rphase = 0;
  int cur = 0;
  TileTotals w;
  const size_t n_chunks = (p.n_spans + JT_CH - 1) / JT_CH;
  // The chunk's ranges (16 B per span, contiguous) come by TMA bulk copy, one chunk ahead: thread 0 asks for the next
  // chunk's while this one is worked on, so the tile bounds never wait for a global load.
  auto fetch_ranges = [&](const size_t c, const int buf) {
    const size_t q = c * JT_CH;
    const uint32_t bytes = (uint32_t)min((size_t)JT_CH, p.n_spans - q) * (uint32_t)sizeof(uint4);
    fence_proxy_async();
    mbar_expect_tx(&sm.bar_rng, bytes);
    bulk_g2s(sm.rngbuf[buf], p.ranges + q, bytes, &sm.bar_rng);
  };
  if (tid == 0 && blockIdx.x < n_chunks) fetch_ranges(blockIdx.x, 0);
  const uint64_t W = (uint64_t)p.window_ns;
  const uint64_t Wi = W < (uint64_t)(500 * kMs) ? W : (uint64_t)(500 * kMs);
  // confidence falls with the tier (1.0, 0.9, 0.8, 0.65), so "confidence >= threshold" is "tier <= cut"
  const int cut = (1.0 >= p.threshold) + (0.9 >= p.threshold) + (0.8 >= p.threshold) + (0.65 >= p.threshold);

  for (size_t chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
    const size_t q0 = chunk * JT_CH;
    const int nq = (int)min((size_t)JT_CH, p.n_spans - q0);
    // ---- stage the chunk's spans: input index, ranges, record (two 16-byte halves per span); clear the hash
    const uint4 *rng = sm.rngbuf[cur];
    if ((int)tid < nq) {
      sm.rng2[tid] = p.ranges2[q0 + tid];
      sm.sidx[tid] = p.span_perm ? p.span_perm[q0 + tid] : (uint32_t)(q0 + tid);
    }
    if ((int)tid < nq) {
      const U8 u = ld_nc_v8(p.spans + sm.sidx[tid]);
      sm.span[tid][0] = u.a;
      sm.span[tid][1] = u.b;
    }
    sm.hkey[tid] = 0;            // JT_THREADS == TM_HASH
    sm.hhead[tid] = TM_END;
    mbar_wait(&sm.bar_rng, rphase);   // this chunk's ranges have landed (asked for one chunk ago)
    rphase ^= 1;
    __syncthreads();
    // Only now, when every thread has seen this phase complete, may the barrier be armed again (a thread still polling
    // while two phases complete would wait for a third).  The other buffer held the previous chunk's ranges, which
    // nobody reads past that chunk's last barrier.
    if (tid == 0 && chunk + gridDim.x < n_chunks) fetch_ranges(chunk + gridDim.x, cur ^ 1);
    cur ^= 1;
    // ---- hash of the chunk's non-zero span trace ids; spans that share an id are chained
    if ((int)tid < nq) {
      const uint32_t tr = sm.span[tid][1].y;
      if (tr != 0) {
        uint32_t slot = trace_slot(tr);
        for (;;) {
          const uint32_t prev = atomicCAS(&sm.hkey[slot], 0u, tr);
          if (prev == 0u || prev == tr) break;
          slot = (slot + 1) & (TM_HASH - 1);
        }
        // push front (16-bit exchange through the containing word is not available: serialise per slot with a CAS loop)
        unsigned short *hp = &sm.hhead[slot];
        unsigned short old = *hp, assumed;
        do {
          assumed = old;
          sm.hnext[tid] = assumed;
          old = atomicCAS(hp, assumed, (unsigned short)tid);
        } while (old != assumed);
      }
    }
    int i0 = 0;
    while (i0 < nq) {
      const uint32_t A = rng[i0].x & ~3u;   // 16-byte aligned start for every column
      if (tid == 0) { sm.cls_min = 0xffffffffu; sm.cls_max = 0u; }   // published by the barrier below
      if (FUSED && (int)tid < JT_CH) { sm.n_note[tid] = 0; sm.n_tr[tid] = 0; sm.n_tr_in[tid] = 0; }   // the loader already notes
      // spans of the chunk whose outer range still ends inside a tile that starts at A (ranges are monotone)
      const bool fits = (int)tid >= i0 && (int)tid < nq && rng[tid].y - A <= (uint32_t)(TM_CAP - 4);
      const int n_fit = __syncthreads_count(fits);
      if (n_fit == 0) {
        // this span's own range is larger than a tile: leave it to k_join_big
        if (tid == 0) big_list[atomicAdd(big_count, 1u)] = (uint32_t)(q0 + i0);
        i0 += 1;
        continue;
      }
      const int i1 = i0 + n_fit;
      const uint32_t Bend = (rng[i1 - 1].y + 3u) & ~3u;
      const uint32_t count = Bend > A ? Bend - A : 0;   // <= TM_CAP, multiple of 4; the columns are padded past n
      if (!FUSED && (int)tid < JT_CH) { sm.n_note[tid] = 0; sm.n_tr[tid] = 0; sm.n_tr_in[tid] = 0; }
      const uint32_t live_beg = rng[i0].x - A, live_end = min(count, rng[i1 - 1].y - A);
      uint32_t cmin = 0xffffffffu, cmax = 0u;
      // tier 1 for one tile event: probe the hash of the chunk's span trace ids, note the spans it falls into
      auto probe_trace = [&](const uint32_t o, const uint32_t tr, const int64_t e_ts) {
        if (tr == 0) return;
        uint32_t slot = trace_slot(tr);
        uint32_t key;
        while ((key = sm.hkey[slot]) != 0u && key != tr) slot = (slot + 1) & (TM_HASH - 1);
        if (key != tr) return;
        for (uint32_t k = sm.hhead[slot]; k != TM_END; k = sm.hnext[k]) {
          if ((int)k < i0 || (int)k >= i1) continue;                 // that span belongs to another tile of the chunk
          const uint4 rg = rng[k];
          if (o + A < rg.x || o + A >= rg.y) continue;                // outside its range: cannot be within the window
          const int64_t s_ts = (int64_t)(((uint64_t)sm.span[k][0].y << 32) | sm.span[k][0].x);
          const uint64_t adt = lsg_absdiff(s_ts, e_ts);
          if (adt > W) continue;                                      // outer window, dns.go:54-56
          atomicAdd(&sm.n_tr[k], 1u);
          if (adt <= Wi) atomicAdd(&sm.n_tr_in[k], 1u);
          if (cut >= 1) {
            const uint32_t at = atomicAdd(&sm.n_note[k], 1u);
            if (at < TM_LIST) sm.note[at][k] = (uint16_t)((LSG_TIER_TRACE << 12) | o);
          }
        }
      };
      if (FUSED) {
        // every thread gathers its share of the tile (at most TM_CAP / JT_THREADS = 7 records): first all its
        // permutation entries (coalesced), then the records four at a time as independent 256-bit loads
        constexpr int PER = (TM_CAP + JT_THREADS - 1) / JT_THREADS;
        uint32_t src[PER];
#pragma unroll
        for (int u = 0; u < PER; ++u) {
          const uint32_t o = tid + u * JT_THREADS;
          src[u] = o < live_end ? __ldg(p.ev_perm + A + o) : 0xffffffffu;
        }
#pragma unroll
        for (int ub = 0; ub < PER; ub += LSG_JT_DEPTH) {
          U8 r[LSG_JT_DEPTH];
#pragma unroll
          for (int u = 0; u < LSG_JT_DEPTH; ++u) {
            r[u].a = make_uint4(0, 0, 0, 0);
            r[u].b = make_uint4(0xffffffffu, 0, 0, 0);   // padding: an identity row no table has
#ifdef LSG_DEBUG_FAKE_GATHER   // timing experiment only (wrong results): every record comes from the same 32 KB, i.e. from L1
            if (ub + u < PER && src[ub + u] != 0xffffffffu) r[u] = ld_nc_v8(p.events_in + (src[ub + u] & 1023u));
#else
            if (ub + u < PER && src[ub + u] != 0xffffffffu) r[u] = ld_nc_v8(p.events_in + src[ub + u]);
#endif
          }
#pragma unroll
          for (int u = 0; u < LSG_JT_DEPTH; ++u) {
            const uint32_t o = tid + (ub + u) * JT_THREADS;
            if (ub + u >= PER || o >= count) continue;
            const uint32_t row = r[u].b.x;
            const uint2 pc = row < p.n_idents ? __ldg(p.ident_pc + row) : make_uint2(0u, 0u);
            sm.ts[o] = (int64_t)(((uint64_t)r[u].a.y << 32) | r[u].a.x);
            sm.pod[o] = pc.x;
            sm.cls[o] = pc.y;
            sm.pid[o] = r[u].b.y;
            sm.conn[o] = r[u].b.z;
            sm.trsig[o] = r[u].b.w;
            if (o >= live_beg && o < live_end) { cmin = min(cmin, pc.y); cmax = max(cmax, pc.y); }
            // tier 1 while the record is still in registers: no second pass over the tile, one barrier less
            if (o < live_end) probe_trace(o, r[u].b.w >> 4, (int64_t)(((uint64_t)r[u].a.y << 32) | r[u].a.x));
          }
        }
      } else if (count) {
        if (tid == 0) {
          fence_proxy_async();
          mbar_expect_tx(&sm.bar, count * 24u);
          bulk_g2s(sm.ts, p.ev.ts + A, count * 8u, &sm.bar);
          bulk_g2s(sm.trsig, p.ev.trsig + A, count * 4u, &sm.bar);
          bulk_g2s(sm.pod, p.ev.ident + A, count * 4u, &sm.bar);
          bulk_g2s(sm.pid, p.ev.pid + A, count * 4u, &sm.bar);
          bulk_g2s(sm.conn, p.ev.conn + A, count * 4u, &sm.bar);
        }
        mbar_wait(&sm.bar, phase);
        phase ^= 1;
        // identity row -> (pod, class), once per staged event instead of once per (span, event) pair; and whether
        // the events the chunk's spans can reach all carry one class (then tier 4 is counted, not enumerated)
        for (uint32_t o = tid; o < count; o += JT_THREADS) {
          const uint32_t row = sm.pod[o];
          const uint2 pc = row < p.n_idents ? __ldg(p.ident_pc + row) : make_uint2(0u, 0u);   // padding past the last event
          sm.pod[o] = pc.x;
          sm.cls[o] = pc.y;
          if (o >= live_beg && o < live_end) { cmin = min(cmin, pc.y); cmax = max(cmax, pc.y); }
        }
      }
      if (count) {
        cmin = __reduce_min_sync(0xffffffffu, cmin);
        cmax = __reduce_max_sync(0xffffffffu, cmax);
        if (lane == 0 && cmin <= cmax) { atomicMin(&sm.cls_min, cmin); atomicMax(&sm.cls_max, cmax); }
      }
      __syncthreads();
      // ---- tier 1 (unfused): every event of the tile probes the hash once
      if (!FUSED) {
        for (uint32_t o = tid; o < live_end; o += JT_THREADS) probe_trace(o, sm.trsig[o] >> 4, sm.ts[o]);
        __syncthreads();   // the hash pass (short, evenly spread) is complete: from here on a span only needs its own two lanes
      }
      // ---- tiers 2-4: thread (2k + h) walks half h of span k's inner range
      const bool uniform = sm.cls_min == sm.cls_max;
      const int k = i0 + (int)(tid >> 1);
      const unsigned h = tid & 1;
      const bool live = k < i1;
      unsigned n_match = 0, n_low = 0, n_unlisted = 0;
      SpanRec s;
      s.ts = 0; s.ident = 0; s.pid = 0; s.conn = 0; s.trace = 0; s.incident = LSG_NONE;
      if (live) {
        s = span_from_halves(sm.span[k][0], sm.span[k][1]);
        const uint4 rg = rng[k];
        // `fanout` trace-id matches (confidence 1.0) already noted: a tier 2-4 candidate can be counted but never kept
        const bool full = cut >= 1 && sm.n_tr[k] >= (uint32_t)p.fanout;
        if (rg.z < rg.w && rg.z != 0xffffffffu) {
          const uint2 spc = __ldg(p.ident_pc + s.ident);
          // an empty pod / service / node never matches (dns.go:61-73); classes are <= n_idents, so ~0 is no event's class
          const bool pod_ok = spc.x != 0;
          const uint32_t s_pod = spc.x, s_cls = spc.y != 0 ? spc.y : 0xffffffffu;
          const bool pid_ok = s.pid != 0, conn_ok = s.conn != 0;
          const uint32_t zo = rg.z - A, wo = rg.w - A;
          PairSpan ps;
          ps.ts = s.ts; ps.s_lo = (uint32_t)(uint64_t)s.ts; ps.pod = spc.x; ps.cls = s_cls; ps.pid = s.pid; ps.conn = s.conn;
          ps.no_trace = s.trace != 0 ? s.trace : 0xffffffffu;   // event trace ids are < 2^28
          ps.pod_ok = pod_ok; ps.pid_ok = pid_ok; ps.conn_ok = conn_ok; ps.full = full;
          const TileCols cols{sm.ts, sm.pod, sm.cls, sm.pid, sm.conn, sm.trsig};
          const auto ident = [](uint32_t o) { return o; };
          if (cut < 4 && uniform) {
            // Every event this chunk can reach carries the same (service, node) class, and a service+node match is
            // below the threshold: it is only counted (correlator.go:83-86).  So count instead of enumerating:
            //   tier 4 = events inside the inner window - trace-id matches among them - tier 2/3 matches,
            // where the window count is the whole quanta [x1, y1) plus an exact test on the two ragged edges, the
            // trace-id matches come from the hash pass, and tiers 2-3 are enumerated on the +-250 ms part only.
            const uint4 r2 = sm.rng2[k];
            const uint32_t x1 = r2.x - A, y1 = r2.y - A, z2 = r2.z - A, w2 = r2.w - A;
            uint32_t n_in = h ? 0u : y1 - x1;
            const uint32_t eb = h ? y1 : zo, ee = h ? wo : x1;      // lane 0: left edge, lane 1: right edge
            const uint32_t *ts_lo = reinterpret_cast<const uint32_t *>(sm.ts);
            for (uint32_t o = eb; o < ee; ++o) {
              if (T32) {
                const int32_t d = (int32_t)(ts_lo[2 * o] - ps.s_lo);
                n_in += (uint32_t)(d < 0 ? -d : d) <= (uint32_t)Wi;
              } else {
                n_in += lsg_absdiff(s.ts, sm.ts[o]) <= Wi;
              }
            }
            const uint32_t m2 = z2 + ((w2 - z2) >> 1);
            const unsigned before = n_match;
            pair_scan<T32, false, TM_LIST>(cols, h ? m2 : z2, h ? w2 : m2, ident, ps, Wi, cut, &sm.n_note[k], &sm.note[0][k], n_match,
                                           n_low, n_unlisted);
            if (s_cls == sm.cls_min) {   // the span belongs to the tile's class: what is left inside the window is tier 4
              uint32_t n4 = n_in - (n_match - before);                             // may wrap per lane; the two-lane sum does not
              if (!h) n4 -= sm.n_tr_in[k];
              n_match += n4;
              n_low += n4;
            }
          } else {
            const uint32_t mid = zo + ((wo - zo) >> 1);
            pair_scan<T32, true, TM_LIST>(cols, h ? mid : zo, h ? wo : mid, ident, ps, Wi, cut, &sm.n_note[k], &sm.note[0][k], n_match,
                                          n_low, n_unlisted);
          }
        }
      }
      __syncwarp();      // the two lanes of a span sit in one warp: their notes are visible to each other now
      // ---- order the span's notes, write its outputs (span_epilogue)
      {
        const TileCols cols{sm.ts, sm.pod, sm.cls, sm.pid, sm.conn, sm.trsig};
        span_epilogue<F, FUSED, TM_LIST>(p, cols, &sm.note[0][live ? k : 0], live ? sm.n_note[k] : 0u, live ? sm.n_tr[k] : 0u, live,
                                         live && rng[k].z == 0xffffffffu, s, live ? (size_t)sm.sidx[k] : (size_t)0, (uint32_t)(q0 + k), h,
                                         lane, cut, n_match, n_low, n_unlisted, [A](uint32_t o) { return A + o; },
                                         [A](uint32_t g) { return g - A; }, big_list, big_count, w);
      }
      __syncthreads();   // every thread is done with the tile before the next bulk copy overwrites it
      i0 = i1;
    }
    __syncthreads();     // the span staging and the hash are reused by the next chunk
  }
  flush_totals(w, sm.cnt, p.counters);
}

#include "join_ring.cuh"

__global__ void k_list_all_spans(uint32_t *__restrict__ big_list, unsigned int *__restrict__ big_count, uint32_t n) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) big_list[i] = i;
  if (blockIdx.x == 0 && threadIdx.x == 0) *big_count = n;
}

// ---- spans whose own range exceeds a tile: one warp per span, straight from global memory ---------------
template <bool FUSED>
__global__ void __launch_bounds__(JT_THREADS)
k_join_big(const __grid_constant__ JoinParams p, const uint32_t *__restrict__ big_list, const unsigned int *__restrict__ big_count) {
  __shared__ Cand s_list[JT_WARPS][JT_LIST];
  __shared__ unsigned long long s_cnt[5];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 5) s_cnt[tid] = 0;
  __syncthreads();
  unsigned long long w_matched = 0, w_low = 0, w_drop = 0, w_cand = 0, w_kept = 0;
  const unsigned n = *big_count;
  const ColsSorted gs{p.ev.ts, p.ev.val, p.ev.ident, p.ev.pid, p.ev.conn, p.ev.trsig, p.ev_perm};
  const ColsIndirect gi{p.events_in, p.ev_perm};
  for (unsigned i = blockIdx.x * JT_WARPS + warp; i < n; i += gridDim.x * JT_WARPS) {
    const size_t q = big_list[i];
    const size_t si = p.span_perm ? p.span_perm[q] : q;
    const SpanRec s = load_span(p.spans + si);
    if (FUSED) process_span(p, gi, s_list[warp], s, si, p.ranges[q], lane, w_matched, w_low, w_drop, w_cand, w_kept);
    else process_span(p, gs, s_list[warp], s, si, p.ranges[q], lane, w_matched, w_low, w_drop, w_cand, w_kept);
  }
  if (lane == 0) {
    if (w_matched) atomicAdd(&s_cnt[0], w_matched);
    if (w_low) atomicAdd(&s_cnt[1], w_low);
    if (w_drop) atomicAdd(&s_cnt[2], w_drop);
    if (w_cand) atomicAdd(&s_cnt[3], w_cand);
    if (w_kept) atomicAdd(&s_cnt[4], w_kept);
  }
  __syncthreads();
  if (tid < 5 && s_cnt[tid]) atomicAdd(&p.counters[tid], s_cnt[tid]);
}

// ---- (pod, class) per identity row -------------------------------------------------------------------
// Rows with the same non-empty (service, node) pair get the same class: 1 + the smallest row index that carries
// the pair, found through an open-addressing table keyed by the pair.
__device__ __forceinline__ uint32_t pair_slot(uint64_t key, uint32_t mask) {
  key ^= key >> 33; key *= 0xff51afd7ed558ccdull; key ^= key >> 33;
  return (uint32_t)key & mask;
}

__global__ void __launch_bounds__(256)
k_ident_class_insert(const lsg_ident *__restrict__ idents, uint32_t n, unsigned long long *__restrict__ tkey,
                     uint32_t *__restrict__ tmin, uint32_t mask) {
  for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
    const lsg_ident id = load_ident(idents, r);
    if (id.service == 0 || id.node == 0) continue;
    const unsigned long long key = ((unsigned long long)id.service << 32) | id.node;   // non-zero
    uint32_t slot = pair_slot(key, mask);
    for (;;) {
      const unsigned long long prev = atomicCAS(&tkey[slot], 0ull, key);
      if (prev == 0ull || prev == key) break;
      slot = (slot + 1) & mask;
    }
    atomicMin(&tmin[slot], r);
  }
}

__global__ void __launch_bounds__(256)
k_ident_class_emit(const lsg_ident *__restrict__ idents, uint32_t n, const unsigned long long *__restrict__ tkey,
                   const uint32_t *__restrict__ tmin, uint32_t mask, uint2 *__restrict__ pc) {
  for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
    const lsg_ident id = load_ident(idents, r);
    uint32_t cls = 0;
    if (id.service != 0 && id.node != 0) {
      const unsigned long long key = ((unsigned long long)id.service << 32) | id.node;
      uint32_t slot = pair_slot(key, mask);
      while (tkey[slot] != key) slot = (slot + 1) & mask;
      cls = tmin[slot] + 1u;
    }
    pc[r] = make_uint2(id.pod, cls);
  }
}

#include "join_slab.cuh"

}  // namespace

int lsg_ident_classes(lsg_ctx *ctx, const lsg_ident *idents, size_t n_idents, const uint2 **out) {
  *out = nullptr;
  if (n_idents >= ((size_t)1 << 31)) return lsg_fail(ctx, LSG_ERR_RANGE, "n_idents %zu is out of range", n_idents);
  size_t cap = 64;
  while (cap < 2 * n_idents) cap <<= 1;
  void *pv, *tv;
  LSG_TRY(lsg_ws(ctx, WS_IDENT_PC, (n_idents + 1) * sizeof(uint2), &pv));
  LSG_TRY(lsg_ws(ctx, WS_IDENT_TABLE, cap * 12, &tv));
  *out = (const uint2 *)pv;
  if (n_idents == 0) return 0;
  unsigned long long *tkey = (unsigned long long *)tv;
  uint32_t *tmin = (uint32_t *)((char *)tv + cap * 8);
  LSG_CUDA(ctx, cudaMemsetAsync(tkey, 0, cap * 8, ctx->stream));
  LSG_CUDA(ctx, cudaMemsetAsync(tmin, 0xff, cap * 4, ctx->stream));
  size_t blocks = (n_idents + 255) / 256, lim = (size_t)ctx->sm_count * 8;
  if (blocks > lim) blocks = lim;
  k_ident_class_insert<<<(unsigned)blocks, 256, 0, ctx->stream>>>(idents, (uint32_t)n_idents, tkey, tmin, (uint32_t)cap - 1);
  LSG_LAUNCH_CHECK(ctx);
  k_ident_class_emit<<<(unsigned)blocks, 256, 0, ctx->stream>>>(idents, (uint32_t)n_idents, tkey, tmin, (uint32_t)cap - 1, (uint2 *)pv);
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}

int lsg_launch_join(lsg_ctx *ctx, const JoinArgs &a) {
  LSG_CUDA(ctx, cudaMemsetAsync(a.counters, 0, 8 * sizeof(uint64_t), ctx->stream));
  if (a.n_spans == 0) return 0;
  void *rv = nullptr;
  LSG_TRY(lsg_ws(ctx, WS_RANGES, a.n_spans * (2 * sizeof(uint4) + sizeof(uint32_t)), &rv));
  uint4 *ranges2 = (uint4 *)rv + a.n_spans;
  const uint32_t n_live = (uint32_t)a.ev.n_live;
  {
    ProfScope ps(ctx, PK_RANGES);
    size_t blocks = (a.n_spans + 255) / 256, cap = (size_t)ctx->sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (a.ev.key.bits <= 31)
      k_span_ranges<uint32_t><<<(unsigned)blocks, 256, 0, ctx->stream>>>(a.spans, a.span_perm, a.n_spans, (const uint32_t *)a.ev.keys,
                                                                      a.ev.bins, n_live, a.ev.key, a.idents, (uint32_t)a.n_idents,
                                                                      a.window_ns, (uint4 *)rv, ranges2, (unsigned long long *)a.counters);
    else
      k_span_ranges<uint64_t><<<(unsigned)blocks, 256, 0, ctx->stream>>>(a.spans, a.span_perm, a.n_spans, (const uint64_t *)a.ev.keys,
                                                                      a.ev.bins, n_live, a.ev.key, a.idents, (uint32_t)a.n_idents,
                                                                      a.window_ns, (uint4 *)rv, ranges2, (unsigned long long *)a.counters);
    LSG_LAUNCH_CHECK(ctx);
  }
  JoinParams p;
  p.spans = a.spans; p.span_perm = a.span_perm; p.n_spans = a.n_spans;
  p.ev = a.ev.col; p.events_in = a.events_in; p.ev_perm = a.ev.perm; p.ranges = (const uint4 *)rv; p.ranges2 = ranges2;
  p.idents = a.idents; p.n_idents = (uint32_t)a.n_idents;
  p.ident_pc = a.ident_pc;
  if (!p.ident_pc) LSG_TRY(lsg_ident_classes(ctx, a.idents, a.n_idents, &p.ident_pc));
  p.window_ns = a.window_ns; p.threshold = a.threshold; p.fanout = a.fanout;
  p.attr_base = a.attr_base; p.cand_event = a.cand_event; p.cand_tier = a.cand_tier; p.attr_out = a.attr_out;
  p.conf_out = a.conf_out; p.retrieval_out = a.retrieval_out; p.n_cand_out = a.n_cand_out;
  p.cand_value = a.cand_value; p.cand_sig = a.cand_sig;
  p.counters = (unsigned long long *)a.counters;
  p.inc_keys = a.inc_keys; p.n_incidents = a.n_incidents;
  p.feat = a.feat;
  static const int seg_env = getenv("LSG_RING_SEG") ? atoi(getenv("LSG_RING_SEG")) : 0;
  p.seg_max = seg_env > 0 ? (uint32_t)seg_env : RG_SEG_MAX;
  const size_t n_chunks = (a.n_spans + JT_CH - 1) / JT_CH;
  const size_t smem = sizeof(TileSmem);
  // spans that do not fit a tile are listed for k_join_big; the list lives behind the ranges
  uint32_t *big_list = (uint32_t *)((char *)rv + a.n_spans * 2 * sizeof(uint4));
  unsigned int *big_count = (unsigned int *)(a.counters + 6);
  ProfScope ps(ctx, PK_JOIN);
  // persistent grid: exactly as many CTAs as are resident at once (SM count x occupancy), each striding over chunks
  auto launch = [&](auto kernel) -> int {
    LSG_CUDA(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 1;
    LSG_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, JT_THREADS, smem));
    if (occ < 1) occ = 1;
    static const bool say = getenv("LSG_DEBUG_OCC") != nullptr;
    if (say) fprintf(stderr, "k_join_tiles: %d threads, %zu B of shared memory, %d CTAs per SM\n", JT_THREADS, smem, occ);
    size_t blocks = n_chunks, cap = (size_t)ctx->sm_count * occ;
    if (blocks > cap) blocks = cap;
    kernel<<<(unsigned)blocks, JT_THREADS, smem, ctx->stream>>>(p, big_list, big_count);
    return 0;
  };
  // inner ranges reach at most min(W, 500 ms) + two time quanta from the span: 32-bit deltas when that is < 2^31 ns
  const int64_t wi = a.window_ns < 500 * kMs ? a.window_ns : 500 * kMs;
  const bool t32 = a.ev.key.shift < 29 && wi + ((int64_t)2 << a.ev.key.shift) < ((int64_t)1 << 31);
  // no sorted copy of the events (lsg_sort_events_finish was told not to gather): the kernels read through the permutation
  const bool fused = a.ev.col.ts == nullptr;
  if (fused && !a.events_in && a.ev.n != 0)
    return lsg_fail(ctx, LSG_ERR_INVALID, "join without sorted columns needs the input records");
  // Which form of the fused tile kernel.  Measured (profiles/r02_summary.md), join ms, fixed chunks / sliding tile:
  // C2 (10 events per span) 0.590 / 0.619, C3 (20 events per span) 66.8 / 52.3.  With few events per span a chunk of
  // 128 spans fits one tile and the halo it gathers again is in L2 already (the CTA next door has just fetched it);
  // with many, every chunk needs two tiles and the halo is most of the second.  So: sliding tile when 128 spans reach
  // more events than a fixed tile holds after its halo (~13 events per span), fixed chunks below.
  // lsg_set_join_path 2 / 3 or LSG_JOIN_RING=1 / 0 force one form.
  static const int ring_force = getenv("LSG_JOIN_RING") ? atoi(getenv("LSG_JOIN_RING")) : -1;
  bool ring_env = a.ev.n >= 13 * a.n_spans;
  if (ring_force >= 0) ring_env = ring_force != 0;
  if (ctx->join_path == 2) ring_env = true;
  if (ctx->join_path == 3) ring_env = false;
  auto launch_ring = [&](auto kernel) -> int {
    const size_t rsmem = sizeof(RingSmem);
    LSG_CUDA(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsmem));
    int occ = 1;
    LSG_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, JT_THREADS, rsmem));
    if (occ < 1) occ = 1;
    static const bool say = getenv("LSG_DEBUG_OCC") != nullptr;
    if (say) fprintf(stderr, "k_join_ring: %d threads, %zu B of shared memory, %d CTAs per SM\n", JT_THREADS, rsmem, occ);
    size_t blocks = n_chunks, cap = (size_t)ctx->sm_count * occ;
    if (blocks > cap) blocks = cap;
    kernel<<<(unsigned)blocks, JT_THREADS, rsmem, ctx->stream>>>(p, big_list, big_count, (unsigned long long *)(a.counters + 7));
    return 0;
  };
  if (fused && ring_env && a.fanout <= TM_MAX_F) {
    if (a.fanout <= 3) { if (t32) LSG_TRY(launch_ring(k_join_ring<3, true>)); else LSG_TRY(launch_ring(k_join_ring<3, false>)); }
    else { if (t32) LSG_TRY(launch_ring(k_join_ring<TM_MAX_F, true>)); else LSG_TRY(launch_ring(k_join_ring<TM_MAX_F, false>)); }
  } else if (a.fanout <= 3) {
    if (fused) { if (t32) LSG_TRY(launch(k_join_tiles<3, true, true>)); else LSG_TRY(launch(k_join_tiles<3, false, true>)); }
    else { if (t32) LSG_TRY(launch(k_join_tiles<3, true, false>)); else LSG_TRY(launch(k_join_tiles<3, false, false>)); }
  } else if (a.fanout <= TM_MAX_F) {
    if (fused) { if (t32) LSG_TRY(launch(k_join_tiles<TM_MAX_F, true, true>)); else LSG_TRY(launch(k_join_tiles<TM_MAX_F, false, true>)); }
    else { if (t32) LSG_TRY(launch(k_join_tiles<TM_MAX_F, true, false>)); else LSG_TRY(launch(k_join_tiles<TM_MAX_F, false, false>)); }
  } else {
    // more kept candidates than the tile kernel holds in registers: every span goes to the warp-per-span kernel
    k_list_all_spans<<<(unsigned)std::min<size_t>((a.n_spans + 255) / 256, (size_t)ctx->sm_count * 8), 256, 0, ctx->stream>>>(
        big_list, big_count, (uint32_t)a.n_spans);
  }
  LSG_LAUNCH_CHECK(ctx);
  {
    size_t blocks = (a.n_spans + JT_WARPS - 1) / JT_WARPS, cap = (size_t)ctx->sm_count * 4;
    if (blocks > cap) blocks = cap;
    if (fused) k_join_big<true><<<(unsigned)blocks, JT_THREADS, 0, ctx->stream>>>(p, big_list, big_count);
    else k_join_big<false><<<(unsigned)blocks, JT_THREADS, 0, ctx->stream>>>(p, big_list, big_count);
  }
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}

// ---- the slab join on bucketed records (join_slab.cuh) -------------------------------------------------------------
bool lsg_bucket_path_ok(const KeySpec &bk, size_t n_events, size_t n_spans, int64_t window_ns, int fanout) {
  (void)window_ns; (void)fanout;
  if (n_events >= ((size_t)1 << 32) || n_spans >= ((size_t)1 << 32)) return false;
  const uint64_t nbk = bk.dead >> LSG_BUCKET_SHIFT;
  if (nbk > ((uint64_t)1 << 24)) return false;                       // tables of at most 64 MB
  if (nbk > 4 * (uint64_t)(n_events + n_spans) + 65536) return false;   // group ids far sparser than the records
  return true;
}

int lsg_launch_join_slab(lsg_ctx *ctx, const JoinArgs &a, const Bucketed &ev, const Bucketed &sp) {
  LSG_CUDA(ctx, cudaMemsetAsync(a.counters, 0, 8 * sizeof(uint64_t), ctx->stream));
  if (a.n_spans == 0) return 0;
  SlabParams p;
  memset(&p, 0, sizeof p);
  JoinParams &j = p.j;
  j.spans = a.spans; j.n_spans = a.n_spans; j.events_in = a.events_in;
  j.idents = a.idents; j.n_idents = (uint32_t)a.n_idents;
  j.ident_pc = a.ident_pc;
  if (!j.ident_pc) LSG_TRY(lsg_ident_classes(ctx, a.idents, a.n_idents, &j.ident_pc));
  j.window_ns = a.window_ns; j.threshold = a.threshold; j.fanout = a.fanout;
  j.attr_base = a.attr_base; j.cand_event = a.cand_event; j.cand_tier = a.cand_tier; j.attr_out = a.attr_out;
  j.conf_out = a.conf_out; j.retrieval_out = a.retrieval_out; j.n_cand_out = a.n_cand_out;
  j.cand_value = a.cand_value; j.cand_sig = a.cand_sig;
  j.counters = (unsigned long long *)a.counters;
  j.inc_keys = a.inc_keys; j.n_incidents = a.n_incidents;
  j.feat = a.feat;
  p.ev = (const lsg_event *)ev.rec; p.ev_orig = ev.orig; p.ev_ends = ev.ends;
  p.sp = (const lsg_span *)sp.rec; p.sp_orig = sp.orig; p.sp_ends = sp.ends;
  p.n_buckets = ev.n_buckets;
  p.ks = ev.key;
  p.gshift = ev.key.tbits - LSG_BUCKET_SHIFT;
  // tile list: cuts of about equal weight, plus the chunks of spans that can match nothing
  const size_t n_ev = a.ev.n;
  const uint32_t n_cuts = (uint32_t)((n_ev + (size_t)JS_SPAN_WEIGHT * a.n_spans) / JS_TILE_TARGET + 2);
  const uint32_t n_dead_chunks = (uint32_t)(a.n_spans / JS_CH + 1);
  void *tv = nullptr, *bv = nullptr;
  LSG_TRY(lsg_ws(ctx, WS_TILES, 64 + ((size_t)n_cuts + n_dead_chunks) * sizeof(uint2), &tv));
  LSG_TRY(lsg_ws(ctx, WS_RANGES, a.n_spans * sizeof(uint32_t) + 64, &bv));
  uint32_t *hdr = (uint32_t *)tv;   // [0] tiles listed
  uint2 *tiles = (uint2 *)((char *)tv + 64);
  p.tiles = tiles; p.n_tiles = hdr;
  uint32_t *big_list = (uint32_t *)bv;
  unsigned int *big_count = (unsigned int *)(a.counters + 6);
  {
    ProfScope ps(ctx, PK_TILES);
    LSG_CUDA(ctx, cudaMemsetAsync(hdr, 0, 64, ctx->stream));
    const uint32_t n_thr = n_cuts + n_dead_chunks;
    k_slab_tiles<<<(n_thr + 255) / 256, 256, 0, ctx->stream>>>(ev.ends, sp.ends, p.n_buckets, n_cuts, tiles, hdr);
    LSG_LAUNCH_CHECK(ctx);
  }
  ProfScope ps(ctx, PK_JOIN);
  const int64_t quantum = (int64_t)1 << ev.key.shift;
  const bool t32 = ev.key.shift < 31 && a.window_ns + quantum < ((int64_t)1 << 31);
  // two CTAs per SM: each takes half of the shared memory an SM can give, the tile capacity follows from it
  int dev_smem = 0, dev_resv = 0;
  LSG_CUDA(ctx, cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, ctx->device));
  LSG_CUDA(ctx, cudaDeviceGetAttribute(&dev_resv, cudaDevAttrReservedSharedMemoryPerBlock, ctx->device));
  const size_t fixed = (sizeof(SlabFixed) + 15) & ~(size_t)15;
  static const int occ_want = getenv("LSG_SLAB_OCC") ? atoi(getenv("LSG_SLAB_OCC")) : 3;   // CTAs per SM (2 or 3): 3 measured 15 % faster on C2
  const int occ_t = occ_want == 2 || a.fanout > 3 ? 2 : 3;
  size_t budget = (size_t)dev_smem / occ_t - (size_t)dev_resv - 256;
  uint32_t cap = (uint32_t)((budget - fixed) / slab_bytes_per_event(t32));
  if (cap > 4088) cap = 4088;   // a note keeps the tile position in 12 bits
  cap &= ~7u;
  p.cap = cap;
  const size_t smem = fixed + (size_t)cap * slab_bytes_per_event(t32) + 16;
  auto launch = [&](auto kernel) -> int {
    LSG_CUDA(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 1;
    LSG_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, JS_THREADS, smem));
    if (occ < 1) occ = 1;
    kernel<<<(unsigned)(ctx->sm_count * occ), JS_THREADS, smem, ctx->stream>>>(p, big_list, big_count);
    return 0;
  };
  if (a.fanout <= 3) {
    if (occ_t == 3) { if (t32) LSG_TRY(launch(k_join_slab<3, true, 3>)); else LSG_TRY(launch(k_join_slab<3, false, 3>)); }
    else { if (t32) LSG_TRY(launch(k_join_slab<3, true, 2>)); else LSG_TRY(launch(k_join_slab<3, false, 2>)); }
  } else if (a.fanout <= TM_MAX_F) {
    if (t32) LSG_TRY(launch(k_join_slab<TM_MAX_F, true, 2>)); else LSG_TRY(launch(k_join_slab<TM_MAX_F, false, 2>));
  } else {
    // more kept candidates than the tile kernel holds in registers: every span goes to the warp-per-span kernel
    k_list_bucketed_spans<<<(unsigned)std::min<size_t>((a.n_spans + 255) / 256, (size_t)ctx->sm_count * 8), 256, 0, ctx->stream>>>(
        big_list, big_count, (uint32_t)a.n_spans);
  }
  LSG_LAUNCH_CHECK(ctx);
  {
    size_t blocks = (a.n_spans + JT_WARPS - 1) / JT_WARPS, capb = (size_t)ctx->sm_count * 4;
    if (blocks > capb) blocks = capb;
    k_join_big_slab<<<(unsigned)blocks, JT_THREADS, 0, ctx->stream>>>(p, big_list, big_count);
  }
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}
