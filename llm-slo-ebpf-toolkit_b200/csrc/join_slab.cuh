// join_slab.cuh — the slab join: Correlator.EnrichAttributes (ebpfcorrelator/correlator.go:50-125) on events and
// spans that the bucket partition (sort.cu) has MOVED into (join group, time slab) buckets.  Included by join.cu
// inside its anonymous namespace (it shares TopF, process_span and the note format with the kernels there).
//
// A bucket is one join group x 64 time quanta, a little longer than the window, so everything a span of bucket b can
// match lies in buckets b-1 .. b+1 of its group.  A tile is a run of consecutive buckets of one group (its "core")
// plus one halo bucket on each side:
//   stage   the tile's events are one contiguous piece of the bucketed array: sequential 256-bit loads.  Inside a
//           bucket the events are in no particular order, so staging is a counting sort by time quantum in shared
//           memory: pass A takes every event's quantum and its rank (shared-memory atomic), a scan turns the counts
//           into quantum starts, pass B reads the records again (L2) and writes the columns at start + rank.
//   ranges  a span's inner window, its +-250 ms part and the quanta wholly inside the window are reads of that
//           quantum table: no global range table, no binary search.
//   tier 1  (dns.go:58-60) hash join as in k_join_tiles: the sub-chunk's span trace ids in a 256-slot hash, every
//           staged event probes it once.
//   tiers 2-4 (dns.go:61-73) one warp per span of a sub-chunk of 128 spans, the lanes stride over the span's pairs:
//           31 of 32 lanes busy whatever the spans' own counts are.  One 128-bit shared load brings (time, pod, pid,
//           conn) of an event; counts are kept per lane and summed once per span, notes are appended at ballot ranks.
//           (Equal per-thread slices of the sub-chunk's pairs laid end to end — merge-path style — were measured
//           first: the nested per-thread loops diverge and ran at 12 of 32 lanes.)
//   order / outputs   as k_join_tiles: notes of two bytes, ordered afterwards in registers under the reference's
//           (confidence desc, |dt| asc, input order), two lanes per span.
// Tiles are cut from the list by every warp on its own (same reads, same result: nothing to exchange), strided over the
// persistent CTAs.
// Spans that do not fit (a bucket with its halos larger than a tile, more notes than the list holds, fanout > 8) go
// to k_join_big_slab: one warp per span over its three buckets straight from global memory.

constexpr int JS_THREADS = 256;   // (512 threads x 2 CTAs measured the same as 256 x 2; 256 x 3 with smaller tiles is the fastest)
constexpr int JS_CH = 128;        // spans per sub-chunk
constexpr int JS_LIST = 32;       // noted candidates per span
constexpr int JS_QCAP = 1024;     // time quanta per tile: 16 buckets
constexpr int JS_MAXB = JS_QCAP >> LSG_BUCKET_SHIFT;
constexpr int JS_HASH = 256;
constexpr uint32_t JS_SPAN_WEIGHT = 10;     // a span weighs this many events when tiles are cut
constexpr uint32_t JS_TILE_TARGET = 2000;   // weight per tile (load balance only: the kernel re-cuts what does not fit)

struct SlabParams {
  JoinParams j;              // parameters and outputs shared with the other kernels (spans / events_in = caller's arrays)
  const lsg_event *ev;       // bucketed events
  const uint32_t *ev_orig;   // bucketed position -> input index
  const uint32_t *ev_ends;   // ends[b]; ends[-1] = 0
  const lsg_span *sp;        // bucketed spans
  const uint32_t *sp_orig;
  const uint32_t *sp_ends;   // n_buckets + 1 entries: the last bucket holds the spans that can match nothing
  uint32_t n_buckets, gshift;   // buckets per group = 1 << gshift
  KeySpec ks;
  const uint2 *tiles;        // (first bucket, end bucket); first == n_buckets: chunk `end` of the dead spans
  const uint32_t *n_tiles;
  uint32_t cap;              // events per tile
};

__device__ __forceinline__ uint32_t bk_start(const uint32_t *ends, uint32_t b) { return __ldg(ends + (int64_t)b - 1); }

// ---- tile list: cut the buckets into runs of about equal weight (events + 10 x spans) ---------------------------
__global__ void __launch_bounds__(256)
k_slab_tiles(const uint32_t *__restrict__ ev_ends, const uint32_t *__restrict__ sp_ends, uint32_t nbk, uint32_t n_cuts,
             uint2 *__restrict__ tiles, uint32_t *__restrict__ n_tiles) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  auto weight_at = [&](uint32_t b) -> uint64_t {   // weight of everything before bucket b
    return (uint64_t)bk_start(ev_ends, b) + (uint64_t)JS_SPAN_WEIGHT * bk_start(sp_ends, b);
  };
  auto first_at_least = [&](uint64_t w) -> uint32_t {
    uint32_t lo = 0, hi = nbk;   // answer in [0, nbk]
    while (lo < hi) {
      const uint32_t mid = lo + ((hi - lo) >> 1);
      if (weight_at(mid) < w) lo = mid + 1; else hi = mid;
    }
    return lo;
  };
  if (t < n_cuts) {
    const uint32_t b0 = first_at_least((uint64_t)t * JS_TILE_TARGET), b1 = first_at_least((uint64_t)(t + 1) * JS_TILE_TARGET);
    if (b0 < b1 && bk_start(sp_ends, b1) > bk_start(sp_ends, b0)) tiles[atomicAdd(n_tiles, 1u)] = make_uint2(b0, b1);
  } else {
    // the spans that can match nothing, 128 at a time: they still get their (empty) outputs
    const uint32_t c = t - n_cuts;
    const uint32_t n_dead = bk_start(sp_ends, nbk + 1) - bk_start(sp_ends, nbk);
    if ((uint64_t)c * JS_CH < n_dead) tiles[atomicAdd(n_tiles, 1u)] = make_uint2(nbk, c);
  }
}

// events in bucket order behind an index array: what process_span needs for the warp-per-span fallback
struct ColsBucketed {
  const lsg_event *ev_;
  const uint32_t *orig_;
  __device__ __forceinline__ int64_t ts(uint32_t o) const { return ev_[o].ts_ns; }
  __device__ __forceinline__ double val(uint32_t o) const { return ev_[o].value; }
  __device__ __forceinline__ uint32_t ident(uint32_t o) const { return ev_[o].ident; }
  __device__ __forceinline__ uint32_t pid(uint32_t o) const { return ev_[o].pid; }
  __device__ __forceinline__ uint32_t conn(uint32_t o) const { return ev_[o].conn; }
  __device__ __forceinline__ uint32_t trsig(uint32_t o) const { return ev_[o].trace_sig; }
  __device__ __forceinline__ uint32_t perm(uint32_t o) const { return orig_[o]; }
  __device__ __forceinline__ uint4 trsig4(uint32_t o, uint32_t end) const {
    uint4 q = make_uint4(0, 0, 0, 0);
    if (o < end) q.x = trsig(o);
    if (o + 1 < end) q.y = trsig(o + 1);
    if (o + 2 < end) q.z = trsig(o + 2);
    if (o + 3 < end) q.w = trsig(o + 3);
    return q;
  }
};

// bucket of a span that can match (the same arithmetic as the partition's SpanBucketFn); n_buckets otherwise
__device__ __forceinline__ uint32_t slab_span_bucket(const SlabParams &p, const SpanRec &s) {
  const uint32_t g = s.ident < p.j.n_idents ? load_ident(p.j.idents, s.ident).grp : 0u;
  if (g != 0 && s.ts != LSG_TS_ZERO && s.ts >= p.ks.ts_min && s.ts <= p.ks.ts_hi) {
    const uint64_t k = lsg_make_key(p.ks, g, s.ts);
    if (k < p.ks.dead) return (uint32_t)(k >> LSG_BUCKET_SHIFT);
  }
  return p.n_buckets;
}

__global__ void __launch_bounds__(JT_THREADS)
k_join_big_slab(const __grid_constant__ SlabParams p, const uint32_t *__restrict__ big_list, const unsigned int *__restrict__ big_count) {
  __shared__ Cand s_list[JT_WARPS][JT_LIST];
  __shared__ unsigned long long s_cnt[5];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 5) s_cnt[tid] = 0;
  __syncthreads();
  unsigned long long w_matched = 0, w_low = 0, w_drop = 0, w_cand = 0, w_kept = 0;
  const unsigned n = *big_count;
  const ColsBucketed cols{p.ev, p.ev_orig};
  for (unsigned i = blockIdx.x * JT_WARPS + warp; i < n; i += gridDim.x * JT_WARPS) {
    const uint32_t pos = big_list[i];
    const SpanRec s = load_span(p.sp + pos);
    const size_t si = p.sp_orig[pos];
    const uint32_t b = slab_span_bucket(p, s);
    uint4 rg = make_uint4(0, 0, 0, 0);
    if (b < p.n_buckets) {
      const uint32_t glo = (b >> p.gshift) << p.gshift, ghi = glo + (1u << p.gshift);
      const uint32_t lo = b > glo ? b - 1 : b, hi = b + 2 < ghi ? b + 2 : ghi;
      rg.x = rg.z = bk_start(p.ev_ends, lo);
      rg.y = rg.w = bk_start(p.ev_ends, hi);
    }
    process_span(p.j, cols, s_list[warp], s, si, rg, lane, w_matched, w_low, w_drop, w_cand, w_kept);
  }
  if (lane == 0) {
    if (w_matched) atomicAdd(&s_cnt[0], w_matched);
    if (w_low) atomicAdd(&s_cnt[1], w_low);
    if (w_drop) atomicAdd(&s_cnt[2], w_drop);
    if (w_cand) atomicAdd(&s_cnt[3], w_cand);
    if (w_kept) atomicAdd(&s_cnt[4], w_kept);
  }
  __syncthreads();
  if (tid < 5 && s_cnt[tid]) atomicAdd(&p.j.counters[tid], s_cnt[tid]);
}

// ---- shared memory of k_join_slab ------------------------------------------------------------------------------
struct SlabFixed {
  uint4 span[JS_CH][2];
  uint32_t sidx[JS_CH];
  uint4 rg[JS_CH];     // x, y: inner window (|dt| <= min(W, 500 ms)) as tile positions; z, w: the quanta wholly inside it
  uint4 rg2[JS_CH];    // x, y: the +-250 ms part (tiers 2-3); z, w: outer window (|dt| <= W)
  uint2 spc[JS_CH];    // the span's (pod, service+node class)
  uint32_t n_note[JS_CH], n_tr[JS_CH], n_tr_in[JS_CH], n_match[JS_CH], n_low[JS_CH], n_edge[JS_CH], n_23[JS_CH], n_unl[JS_CH];
  uint16_t note[JS_LIST][JS_CH];
  uint32_t hkey[JS_HASH];
  uint16_t hhead[JS_HASH];
  uint16_t hnext[JS_CH];
  uint32_t qs[JS_QCAP + 8];              // quantum starts (counts while they are taken)
  uint32_t scan_w[JS_THREADS / 32];
  uint32_t cls_min, cls_max;
  unsigned long long cnt[5];
};
constexpr uint32_t JS_UNIT_DONE = 0, JS_UNIT_TILE = 1, JS_UNIT_BIG = 2, JS_UNIT_DEAD = 3;

__host__ __device__ inline uint32_t slab_bytes_per_event(bool t32) { return t32 ? 30u : 34u; }

// The persistent CTA's work loop.  F: kept candidates held in registers; T32: every pair the kernel forms is closer
// than 2^31 ns, so the low words of the times give exact differences.
template <int F, bool T32, int OCC>
__global__ void __launch_bounds__(JS_THREADS, OCC)
k_join_slab(const __grid_constant__ SlabParams p, uint32_t *__restrict__ big_list, unsigned int *__restrict__ big_count) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SlabFixed &sm = *reinterpret_cast<SlabFixed *>(smem_raw);
  const uint32_t cap = p.cap;
  uint4 *q4 = reinterpret_cast<uint4 *>(smem_raw + ((sizeof(SlabFixed) + 15) & ~(size_t)15));   // (ts low word, pod, pid, conn)
  uint32_t *trs = reinterpret_cast<uint32_t *>(q4 + cap);
  uint32_t *cls = trs + cap;
  uint32_t *aux = cls + cap;             // pass A -> pass B: (quantum << 16) | rank
  uint32_t *ts_hi = aux + cap;           // only when !T32
  uint16_t *slot = reinterpret_cast<uint16_t *>(T32 ? ts_hi : ts_hi + cap);   // staged position -> position in the tile's piece
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  const JoinParams &jp = p.j;
  if (tid < 5) sm.cnt[tid] = 0;
  __syncthreads();
  unsigned long long w_matched = 0, w_low = 0, w_drop = 0, w_cand = 0, w_kept = 0;
  const uint64_t W = (uint64_t)jp.window_ns;
  const uint64_t Wi = W < (uint64_t)(500 * kMs) ? W : (uint64_t)(500 * kMs);
  const uint64_t W2 = Wi < (uint64_t)(250 * kMs) ? Wi : (uint64_t)(250 * kMs);
  const int cut = (1.0 >= jp.threshold) + (0.9 >= jp.threshold) + (0.8 >= jp.threshold) + (0.65 >= jp.threshold);
  const uint32_t gmask = (1u << p.gshift) - 1u, nbk = p.n_buckets;
  const uint32_t shift = p.ks.shift;
  const uint32_t n_tiles = *p.n_tiles;
  uint32_t t_cur = blockIdx.x, bc_next = 0, b1_cur = 0;   // the CTA's tile (strided over the list) and the cursor inside it
  bool have_tile = false;

  for (;;) {
    // ---- next unit: the current tile (or the next one) is cut into what fits shared memory.  Every warp does the same
    //      cut from the same (L2-resident) bucket tables, so nothing has to be exchanged and nobody waits for warp 0.
    uint32_t kind = JS_UNIT_DONE, bc = 0, be = 0, lo_b = 0, hi_b = 0, S0 = 0, nS = 0;
    for (;;) {
      if (!have_tile) {
        if (t_cur >= n_tiles) break;
        const uint2 tb = __ldg(p.tiles + t_cur);
        t_cur += gridDim.x;
        if (tb.x == nbk) {   // a chunk of the spans that can match nothing
          kind = JS_UNIT_DEAD;
          S0 = bk_start(p.sp_ends, nbk) + tb.y * JS_CH;
          nS = min((uint32_t)JS_CH, bk_start(p.sp_ends, nbk + 1) - S0);
          bc = be = lo_b = hi_b = nbk;
          break;
        }
        bc_next = tb.x; b1_cur = tb.y; have_tile = true;
      }
      // first bucket at or after bc_next that holds spans
      uint32_t b = bc_next;
      while (b < b1_cur) {
        const uint32_t bb = b + lane;
        const bool has = bb < b1_cur && bk_start(p.sp_ends, bb + 1) > bk_start(p.sp_ends, bb);
        const unsigned m = __ballot_sync(0xffffffffu, has);
        if (m) { b += __ffs(m) - 1; break; }
        b += 32;
      }
      if (b >= b1_cur) { have_tile = false; continue; }
      bc = b;
      const uint32_t glo = bc & ~gmask, ghi = glo + gmask + 1u;
      const uint32_t gend = b1_cur < ghi ? b1_cur : ghi;
      lo_b = bc > glo ? bc - 1 : bc;
      // lane m proposes the core [bc, bc + m + 1): it fits when the loaded buckets [lo, hi) are at most JS_MAXB and
      // hold at most `cap` events
      const uint32_t e_lo = bk_start(p.ev_ends, lo_b);
      const uint32_t cand_be = bc + lane + 1;
      bool fits = false;
      if (cand_be <= gend) {
        const uint32_t h = cand_be + 1 < ghi ? cand_be + 1 : ghi;
        fits = h - lo_b <= (uint32_t)JS_MAXB && bk_start(p.ev_ends, h) - e_lo <= cap;
      }
      const unsigned fm = __ballot_sync(0xffffffffu, fits);
      const int n_fit = __ffs(~fm) - 1;   // conditions are monotone: the leading run of ones (32 lanes > JS_MAXB)
      if (n_fit <= 0) {
        kind = JS_UNIT_BIG;   // this bucket with its halos is larger than a tile
        be = bc + 1;
      } else {
        kind = JS_UNIT_TILE;
        be = bc + (uint32_t)n_fit;
      }
      hi_b = be + 1 < ghi ? be + 1 : ghi;
      bc_next = be;
      S0 = bk_start(p.sp_ends, bc);
      nS = bk_start(p.sp_ends, be) - S0;
      break;
    }
    if (kind == JS_UNIT_DONE) break;
    if (kind == JS_UNIT_BIG) {
      // every span of the bucket goes to the warp-per-span kernel
      for (uint32_t k = tid; k < nS; k += JS_THREADS) big_list[atomicAdd(big_count, 1u)] = S0 + k;
      continue;
    }
    const uint32_t E0 = kind == JS_UNIT_TILE ? bk_start(p.ev_ends, lo_b) : 0u;
    const uint32_t nE = kind == JS_UNIT_TILE ? bk_start(p.ev_ends, hi_b) - E0 : 0u;
    const uint32_t nq = (hi_b - lo_b) << LSG_BUCKET_SHIFT;             // time quanta of the loaded buckets (0 for dead spans)
    const uint64_t q0 = (uint64_t)(lo_b & gmask) << LSG_BUCKET_SHIFT;   // the tile's first quantum on the group's time axis
    __syncthreads();   // the previous unit's last sub-chunk is done with the tables that are rebuilt now
    if (tid == 0) { sm.cls_min = 0xffffffffu; sm.cls_max = 0u; }

    // ---- stage A: quantum and rank of every event
    for (uint32_t q = tid; q < nq + 1; q += JS_THREADS) sm.qs[q] = 0;
    __syncthreads();
    for (uint32_t i = tid; i < nE; i += JS_THREADS) {
      const int64_t ts = __ldg(&p.ev[E0 + i].ts_ns);
      uint32_t q = (uint32_t)((((uint64_t)ts - (uint64_t)p.ks.ts_min) >> shift) - q0);
      if (q >= nq) q = nq - 1;   // cannot happen for a bucketed event; keeps a corrupt input inside the table
      const uint32_t r = atomicAdd(&sm.qs[q], 1u);
      aux[i] = (q << 16) | r;
    }
    __syncthreads();
    // ---- exclusive scan of the quantum counts (four per thread, 1024 + 1 entries at most)
    {
      const uint32_t base = tid * 4;
      uint32_t v[4], t = 0;
#pragma unroll
      for (int u = 0; u < 4; ++u) { v[u] = base + u < nq ? sm.qs[base + u] : 0u; }
#pragma unroll
      for (int u = 0; u < 4; ++u) { const uint32_t c = v[u]; v[u] = t; t += c; }
      uint32_t x = t;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= (unsigned)o) x += y;
      }
      if (lane == 31) sm.scan_w[warp] = x;
      __syncthreads();
      uint32_t pre = x - t;
      for (unsigned w = 0; w < warp; ++w) pre += sm.scan_w[w];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (base + u < nq) sm.qs[base + u] = pre + v[u];
      if (tid == 0) sm.qs[nq] = nE;
    }
    __syncthreads();
    // ---- stage B: the records into the columns at quantum start + rank
    {
      uint32_t cmin = 0xffffffffu, cmax = 0u;
      for (uint32_t i0 = tid; i0 < nE; i0 += 4 * JS_THREADS) {
        U8 r[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t i = i0 + u * JS_THREADS;
          if (i < nE) r[u] = ld_nc_v8(p.ev + E0 + i);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t i = i0 + u * JS_THREADS;
          if (i >= nE) continue;
          const uint32_t a = aux[i];
          const uint32_t pos = sm.qs[a >> 16] + (a & 0xffffu);
          const uint32_t row = r[u].b.x;
          const uint2 pc = row < jp.n_idents ? __ldg(jp.ident_pc + row) : make_uint2(0u, 0u);
          q4[pos] = make_uint4(r[u].a.x, pc.x, r[u].b.y, r[u].b.z);
          if (!T32) ts_hi[pos] = r[u].a.y;
          trs[pos] = r[u].b.w;
          cls[pos] = pc.y;
          slot[pos] = (uint16_t)i;
          cmin = min(cmin, pc.y); cmax = max(cmax, pc.y);
        }
      }
      cmin = __reduce_min_sync(0xffffffffu, cmin);
      cmax = __reduce_max_sync(0xffffffffu, cmax);
      if (lane == 0 && cmin <= cmax) { atomicMin(&sm.cls_min, cmin); atomicMax(&sm.cls_max, cmax); }
    }
    // (the barrier that publishes the columns is the first one of the sub-chunk loop)

    for (uint32_t c0 = 0; c0 < nS; c0 += JS_CH) {
      const int ns = (int)min((uint32_t)JS_CH, nS - c0);
      __syncthreads();   // the columns are complete / the previous sub-chunk is done with the span staging
      // ---- the sub-chunk's spans, their windows as positions of the tile, the trace-id hash
      if (tid < JS_HASH) {
        sm.hkey[tid] = 0;
        sm.hhead[tid] = TM_END;
      }
      if ((int)tid < ns) {
        const U8 u = ld_nc_v8(p.sp + S0 + c0 + tid);
        sm.span[tid][0] = u.a;
        sm.span[tid][1] = u.b;
        sm.sidx[tid] = __ldg(p.sp_orig + S0 + c0 + tid);
        sm.n_note[tid] = 0; sm.n_tr[tid] = 0; sm.n_tr_in[tid] = 0; sm.n_match[tid] = 0; sm.n_low[tid] = 0;
        sm.n_edge[tid] = 0; sm.n_23[tid] = 0; sm.n_unl[tid] = 0;
        uint32_t ox = 0, oy = 0, zo = 0, wo = 0, x1 = 0, y1 = 0, z2 = 0, w2 = 0;
        if (nq) {
          const int64_t s_ts = (int64_t)(((uint64_t)u.a.y << 32) | u.a.x);
          const uint64_t off = (uint64_t)s_ts - (uint64_t)p.ks.ts_min;   // a span that can match lies on the key's time axis
          const uint64_t quantum = (uint64_t)1 << shift;
          auto pos_of = [&](uint64_t qv) -> uint32_t {   // first staged event at or after quantum qv of the group's axis
            if (qv <= q0) return 0u;
            const uint64_t rel = qv - q0;
            return rel >= nq ? nE : sm.qs[rel];
          };
          // NOTE: sm.qs is complete: the barrier at the top of the loop follows the scan
          ox = pos_of(off > W ? (off - W) >> shift : 0);
          oy = pos_of(((off + W) >> shift) + 1);
          zo = pos_of(off > Wi ? (off - Wi) >> shift : 0);
          wo = pos_of(((off + Wi) >> shift) + 1);
          const uint64_t qa = off > Wi ? (off - Wi + quantum - 1) >> shift : 0;   // first quantum starting at or after ts - Wi
          const uint64_t qb1 = (off + Wi + 1) >> shift;                             // first quantum not ending by ts + Wi
          z2 = max(zo, min(wo, pos_of(off > W2 ? (off - W2) >> shift : 0)));
          w2 = max(z2, min(wo, pos_of(((off + W2) >> shift) + 1)));
          if (qa < qb1) {
            x1 = max(zo, min(wo, pos_of(qa)));
            y1 = max(x1, min(wo, pos_of(qb1)));
          } else {
            x1 = y1 = z2;   // no whole quantum inside the window (a quantum wider than it): everything is enumerated
          }
        }
        sm.rg[tid] = make_uint4(zo, wo, x1, y1);
        sm.rg2[tid] = make_uint4(z2, w2, ox, oy);
        sm.spc[tid] = nq && u.a.z < jp.n_idents ? __ldg(jp.ident_pc + u.a.z) : make_uint2(0u, 0u);
      }
      __syncthreads();
      const bool uniform = sm.cls_min == sm.cls_max;
      const bool counted = cut < 4 && uniform;   // the service+node tier is counted, not enumerated
      if ((int)tid < ns) {
        const uint32_t tr = sm.span[tid][1].y;
        if (tr != 0) {
          uint32_t hs = trace_slot(tr);
          for (;;) {
            const uint32_t prev = atomicCAS(&sm.hkey[hs], 0u, tr);
            if (prev == 0u || prev == tr) break;
            hs = (hs + 1) & (JS_HASH - 1);
          }
          unsigned short *hp = &sm.hhead[hs];
          unsigned short old = *hp, assumed;
          do {
            assumed = old;
            sm.hnext[tid] = assumed;
            old = atomicCAS(hp, assumed, (unsigned short)tid);
          } while (old != assumed);
        }
      }
      __syncthreads();
      // ---- tier 1: every staged event probes the hash once
      for (uint32_t o = tid; o < nE; o += JS_THREADS) {
        const uint32_t tr = trs[o] >> 4;
        if (tr == 0) continue;
        uint32_t hs = trace_slot(tr);
        uint32_t key;
        while ((key = sm.hkey[hs]) != 0u && key != tr) hs = (hs + 1) & (JS_HASH - 1);
        if (key != tr) continue;
        for (uint32_t k = sm.hhead[hs]; k != TM_END; k = sm.hnext[k]) {
          if (o < sm.rg2[k].z || o >= sm.rg2[k].w) continue;             // outside the outer window's quanta
          uint64_t adt;
          if (T32) {   // inside those quanta |dt| < W + quantum < 2^31
            const int32_t d = (int32_t)(q4[o].x - sm.span[k][0].x);
            adt = (uint64_t)(uint32_t)(d < 0 ? -d : d);
          } else {
            const int64_t s_ts = (int64_t)(((uint64_t)sm.span[k][0].y << 32) | sm.span[k][0].x);
            adt = lsg_absdiff(s_ts, (int64_t)(((uint64_t)ts_hi[o] << 32) | q4[o].x));
          }
          if (adt > W) continue;                                        // outer window, dns.go:54-56
          atomicAdd(&sm.n_tr[k], 1u);
          if (adt <= Wi) atomicAdd(&sm.n_tr_in[k], 1u);
          if (cut >= 1) {
            const uint32_t at = atomicAdd(&sm.n_note[k], 1u);
            if (at < JS_LIST) sm.note[at][k] = (uint16_t)((LSG_TIER_TRACE << 12) | o);
          }
        }
      }
      __syncthreads();
      // ---- tiers 2-4: one warp per span, the lanes stride over the span's pairs.  Span data is warp-uniform (a few
      //      broadcast loads of what the staging pass prepared), counts are kept per lane and summed once per span,
      //      notes are appended at ballot ranks (no atomics).
      {
        const uint32_t Wi32 = (uint32_t)Wi;
        for (int k = (int)warp; k < ns; k += JS_THREADS / 32) {
          const uint4 rg = sm.rg[k];
          const uint32_t zo = rg.x, wo = rg.y, x1 = rg.z, y1 = rg.w;
          if (zo >= wo) continue;   // nothing inside the inner window (also: every span that can match nothing)
          const uint4 sa = sm.span[k][0], sb = sm.span[k][1];
          const uint2 spc = sm.spc[k];
          const uint32_t s_lo = sa.x, s_pid = sa.w, s_conn = sb.x;
          const int64_t s_ts = (int64_t)(((uint64_t)sa.y << 32) | sa.x);
          const bool ok2 = spc.x != 0 && s_pid != 0, ok3 = spc.x != 0 && s_conn != 0;   // dns.go:61-68: non-empty pod, pid / conn
          const uint32_t s_cls = spc.y != 0 ? spc.y : 0xffffffffu;
          const uint32_t no_trace = sb.y != 0 ? sb.y : 0xffffffffu;
          const bool full = cut >= 1 && sm.n_tr[k] >= (uint32_t)jp.fanout;   // the kept list is already full of trace-id matches
          // enumerated: everything but the quanta wholly inside the window that lie outside +-250 ms.  Pair j of the
          // span sits at zo + j, shifted by the gaps it has passed.
          uint32_t lenA = wo - zo, lenAB = lenA, total = lenA, gap1 = 0, gap2 = 0;
          if (counted) {
            const uint32_t z2 = sm.rg2[k].x, w2 = sm.rg2[k].y;
            const uint32_t a1 = min(x1, z2), c0s = max(y1, w2);
            lenA = a1 - zo; lenAB = lenA + (w2 - z2); total = lenAB + (wo - c0s);
            gap1 = z2 - a1; gap2 = c0s - w2;
          }
          uint32_t n_note = sm.n_note[k];          // the probe's trace-id notes come first (warp-uniform)
          uint32_t c_match = 0, c_low = 0, c_edge = 0, n_unl = 0;   // per lane, except n_unl
          for (uint32_t jb = 0; jb < total; jb += 32) {
            const uint32_t j = min(jb + lane, total - 1);
            const bool valid = jb + lane < total;
            const uint32_t o = zo + j + (j >= lenA ? gap1 : 0u) + (j >= lenAB ? gap2 : 0u);
            const uint4 e = q4[o];
            bool in, in100, in250;
            if (T32) {
              const int32_t d = (int32_t)(e.x - s_lo);
              const uint32_t ad = (uint32_t)(d < 0 ? -d : d);
              in = ad <= Wi32; in100 = ad <= (uint32_t)(100 * kMs); in250 = ad <= (uint32_t)(250 * kMs);
            } else {
              const uint64_t ad = lsg_absdiff(s_ts, (int64_t)(((uint64_t)ts_hi[o] << 32) | e.x));
              in = ad <= Wi; in100 = ad <= (uint64_t)(100 * kMs); in250 = ad <= (uint64_t)(250 * kMs);
            }
            in &= valid;
            const bool same_pod = e.y == spc.x;
            const bool t2 = ok2 & same_pod & (e.z == s_pid) & in100;        // dns.go:61-64
            const bool t3 = ok3 & same_pod & (e.w == s_conn) & in250;      // dns.go:65-68
            int tier;
            bool m;
            if (counted) {
              tier = t2 ? LSG_TIER_POD_PID : LSG_TIER_POD_CONN;
              m = in & (t2 | t3);
              c_edge += in & ((o < x1) | (o >= y1));
            } else {
              const bool same_sn = cls[o] == s_cls;
              tier = t2 ? LSG_TIER_POD_PID : t3 ? LSG_TIER_POD_CONN : same_sn ? LSG_TIER_SVC_NODE : 0;   // dns.go:69-73
              m = in & (tier != 0);
            }
            if (m) m = (trs[o] >> 4) != no_trace;   // a trace-id match belongs to tier 1 (the hash pass)
            c_match += m;
            const bool low = m & (tier > cut);                                 // correlator.go:83-86
            c_low += low;
            const unsigned cm = __ballot_sync(0xffffffffu, m & !low);          // candidates
            if (cm) {
              if (full) {
                n_unl += __popc(cm);
              } else {
                if (m & !low) {
                  const uint32_t at = n_note + __popc(cm & ((1u << lane) - 1u));
                  if (at < JS_LIST) sm.note[at][k] = (uint16_t)((tier << 12) | o);
                }
                n_note += __popc(cm);
              }
            }
          }
          c_match = __reduce_add_sync(0xffffffffu, c_match);
          c_low = __reduce_add_sync(0xffffffffu, c_low);
          c_edge = __reduce_add_sync(0xffffffffu, c_edge);
          if (lane == 0) {
            sm.n_note[k] = n_note; sm.n_match[k] = c_match; sm.n_low[k] = c_low; sm.n_edge[k] = c_edge;
            sm.n_23[k] = counted ? c_match : 0u;   // in the counted form every enumerated match is a tier 2-3 match
            sm.n_unl[k] = n_unl;
          }
        }
      }
      __syncthreads();
      // ---- order and write: two lanes per span (lane h takes the notes of parity h, the odd lane hands its best F over)
      {
        const int k = (int)(tid >> 1);
        const unsigned h = tid & 1;
        const bool live = k < ns;
        SpanRec s;
        s.ts = 0; s.ident = 0; s.pid = 0; s.conn = 0; s.trace = 0; s.incident = LSG_NONE;
        TopF<F> top;
        top.init();
        uint32_t n_cand = 0;
        bool redo = false;
        if (live) {
          s = span_from_halves(sm.span[k][0], sm.span[k][1]);
          n_cand = sm.n_note[k];
          redo = n_cand > (uint32_t)JS_LIST;
          if (!redo) {
            // four notes at a time: their input indices (the tie-break of the reference's stable sort) are global loads,
            // issued together so that their latency is paid once per batch
            for (uint32_t i = h; i < n_cand; i += 8) {
              uint32_t ent[4], orig[4];
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                const uint32_t ix = i + 2 * u;
                ent[u] = ix < n_cand ? (uint32_t)sm.note[ix][k] : 0xffffffffu;
                orig[u] = ent[u] != 0xffffffffu ? __ldg(p.ev_orig + E0 + slot[ent[u] & 0xfffu]) : 0u;
              }
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                if (ent[u] == 0xffffffffu) continue;
                const uint32_t o = ent[u] & 0xfffu, tier = ent[u] >> 12;
                uint64_t adt;
                if (T32) {   // a noted candidate is within the window of its span
                  const int32_t d = (int32_t)(q4[o].x - (uint32_t)(uint64_t)s.ts);
                  adt = (uint64_t)(uint32_t)(d < 0 ? -d : d);
                } else {
                  adt = lsg_absdiff(s.ts, (int64_t)(((uint64_t)ts_hi[o] << 32) | q4[o].x));
                }
                top.insert(((uint64_t)tier << 56) | adt, orig[u], o);
              }
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
        if (live) {
          if (redo) {
            if (!h) big_list[atomicAdd(big_count, 1u)] = S0 + c0 + (uint32_t)k;   // more candidates than the list holds
          } else {
            const unsigned pair = 3u << (lane & 30u);
            const size_t si = sm.sidx[k];
            double vals[F];
            unsigned sigs[F];
            unsigned present = 0;
            int kept = 0;
#pragma unroll
            for (int r = 0; r < F; ++r) {
              const bool mine = r < jp.fanout && top.idx[r] != 0xffffffffu;
              const uint32_t pos = __shfl_sync(pair, mine ? top.pos[r] : 0xffffffffu, lane & 30u);   // the even lane's final list
              const bool has = pos != 0xffffffffu;
              vals[r] = has ? __ldg(&p.ev[E0 + slot[pos]].value) : 0.0;
              sigs[r] = has ? (trs[pos] & 15u) : 0xffu;
              kept += has;
              present |= sigs[r] < LSG_NSIG ? 1u << sigs[r] : 0u;
              if (!h && r < jp.fanout) {
                jp.cand_event[si * (size_t)jp.fanout + r] = has ? top.idx[r] : LSG_NONE;      // correlator.go:101-104
                jp.cand_tier[si * (size_t)jp.fanout + r] = has ? (uint8_t)(top.k1[r] >> 56) : (uint8_t)0;
                if (jp.cand_value) jp.cand_value[si * (size_t)jp.fanout + r] = vals[r];
                if (jp.cand_sig) jp.cand_sig[si * (size_t)jp.fanout + r] = (uint8_t)sigs[r];
              }
            }
            // per-attribute max over the kept candidates (correlator.go:106-112), six columns per lane
            double v0 = nan, v3 = nan, v4 = nan;
            const bool to_incident = jp.inc_keys && s.incident != LSG_NONE && s.incident < jp.n_incidents;
#pragma unroll
            for (int c = 0; c < LSG_NSIG / 2; ++c) {
              const unsigned col = h * (LSG_NSIG / 2) + c;
              double v = jp.attr_base ? jp.attr_base[si * LSG_NSIG + col] : nan;
              if ((present >> col) & 1u) {
#pragma unroll
                for (int r = 0; r < F; ++r)
                  if (sigs[r] == col && (isnan(v) || vals[r] > v)) v = vals[r];
              }
              if (jp.attr_out) jp.attr_out[si * LSG_NSIG + col] = v;
              if (to_incident && !isnan(v)) {
                atomicMax(&jp.inc_keys[(size_t)s.incident * LSG_NSIG + col], f64_key(v));   // A7 glue: max per incident and signal
                if (jp.feat) feat_add(jp.feat, (size_t)s.incident * LSG_NSIG + col, v);
              }
              if (c == 0) v0 = v;   // on the even lane: dns (0), connect (3), tls (4)
              if (c == 3) v3 = v;
              if (c == 4) v4 = v;
            }
            if (!h) {
              const uint32_t n_tr = sm.n_tr[k];
              uint32_t n_match = sm.n_match[k] + n_tr, n_low = sm.n_low[k];
              if (cut < 1) n_low += n_tr;
              if (counted) {
                // what is left inside the exact inner window after the trace-id and pod matches is the service+node tier
                // (below the threshold here: counted as matched and low-confidence, correlator.go:83-86)
                const uint2 spc = sm.spc[k];
                if (spc.y != 0 && spc.y == sm.cls_min) {
                  const uint32_t n_in = (sm.rg[k].w - sm.rg[k].z) + sm.n_edge[k];
                  const uint32_t n4 = n_in - sm.n_23[k] - sm.n_tr_in[k];
                  n_match += n4;
                  n_low += n4;
                }
              }
              n_cand += sm.n_unl[k];   // candidates that were counted without being noted
              if (jp.conf_out) jp.conf_out[si] = kept ? lsg_tier_conf((int)(top.k1[0] >> 56)) : 0.0;   // correlator.go:113-118
              if (jp.retrieval_out) {   // DecomposeRetrieval, correlator.go:179-194: dns + connect + tls in that order
                double total = 0.0;
                if (!isnan(v0)) total += v0;
                if (!isnan(v3)) total += v3;
                if (!isnan(v4)) total += v4;
                jp.retrieval_out[si] = total;
              }
              if (jp.n_cand_out) jp.n_cand_out[si] = n_cand;
              w_matched += n_match;
              w_low += n_low;
              w_cand += n_cand;
              w_kept += kept;
              if (n_cand > (uint32_t)jp.fanout) w_drop += n_cand - jp.fanout;   // correlator.go:101-103
            }
          }
        }
      }
    }   // sub-chunks
  }     // units
  // counters: lanes -> warp -> block -> one atomic each
  for (int o = 16; o; o >>= 1) {
    w_matched += __shfl_xor_sync(0xffffffffu, w_matched, o);
    w_low += __shfl_xor_sync(0xffffffffu, w_low, o);
    w_drop += __shfl_xor_sync(0xffffffffu, w_drop, o);
    w_cand += __shfl_xor_sync(0xffffffffu, w_cand, o);
    w_kept += __shfl_xor_sync(0xffffffffu, w_kept, o);
  }
  if (lane == 0) {
    if (w_matched) atomicAdd(&sm.cnt[0], w_matched);
    if (w_low) atomicAdd(&sm.cnt[1], w_low);
    if (w_drop) atomicAdd(&sm.cnt[2], w_drop);
    if (w_cand) atomicAdd(&sm.cnt[3], w_cand);
    if (w_kept) atomicAdd(&sm.cnt[4], w_kept);
  }
  __syncthreads();
  if (tid < 5 && sm.cnt[tid]) atomicAdd(&jp.counters[tid], sm.cnt[tid]);
}

__global__ void k_list_bucketed_spans(uint32_t *__restrict__ big_list, unsigned int *__restrict__ big_count, uint32_t n) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) big_list[i] = i;
  if (blockIdx.x == 0 && threadIdx.x == 0) *big_count = n;
}
