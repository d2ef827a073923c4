// join_ring.cuh — the tile kernel of join.cu with a SLIDING tile (included by join.cu; FUSED form only: the events are
// gathered through the permutation).
//
// k_join_tiles cuts the spans into fixed chunks of 128 and gives every chunk to whichever CTA comes next, so each tile
// starts from nothing: it gathers its spans' whole outer range, of which the +-window halo (C2: ~27 % of a tile, C3:
// ~43 %) was already gathered by the CTA that took the chunk before, and a chunk that does not fit one tile leaves a
// second, nearly empty tile behind.  Measured (profiles/r02_summary.md): the kernel's time follows the number of records
// gathered and staged, not DRAM traffic or occupancy.  Here a CTA owns a SEGMENT of consecutive spans (taken from a
// shared cursor, a shrinking share of what is left so that the CTAs finish together) and walks it with one tile whose
// columns are a ring: sorted position g lives in slot g mod RG_CAP.  A tile takes as many spans as fit, up to 128;
// the next tile starts where this one ended and only gathers the positions behind what is already resident.  Events
// that stay are probed again against the new tile's trace-id hash from shared memory (no global access).
//
// Everything else is shared with k_join_tiles: the hash probe for tier 1 (dns.go:58-60) has the same shape, tiers 2-4
// (dns.go:61-73, two threads per span, counted tier 4 in tiles of one (service, node) class) go through pair_scan, and
// the ordering of the notes and the write of the candidates, attributes, confidence and counters
// (correlator.go:92-118) through span_epilogue, with `position mod RG_CAP` where the fixed tile uses `position - A`.
#pragma once

constexpr int RG_CAP = 2048;                 // events resident per CTA (28 B each)
constexpr uint32_t RG_MASK = RG_CAP - 1;
constexpr int RG_LIST = 32;                  // noted candidates per span (a span with more goes to k_join_big)
constexpr int RG_PER = RG_CAP / JT_THREADS;  // records a thread gathers for a tile that starts from nothing
constexpr uint32_t RG_SEG_MAX = 4096;        // spans per segment at most
static_assert((RG_CAP & (RG_CAP - 1)) == 0 && RG_CAP <= 4096, "slot = position & mask, and a note keeps the slot in 12 bits");
static_assert(RG_PER % LSG_JT_DEPTH == 0, "the loader's batches divide a thread's share");

struct RingSmem {
  int64_t ts[RG_CAP];
  uint32_t pod[RG_CAP];             // the identity row's pod id
  uint32_t cls[RG_CAP];             // the row's (service, node) class
  uint32_t pid[RG_CAP];
  uint32_t conn[RG_CAP];
  uint32_t trsig[RG_CAP];
  uint4 span[JT_CH][2];
  uint4 rng[JT_CH];
  uint32_t sidx[JT_CH];
  uint16_t note[RG_LIST][JT_CH];    // entry-major: consecutive spans, consecutive halfwords
  uint32_t n_note[JT_CH];
  uint32_t n_tr[JT_CH];
  uint32_t n_tr_in[JT_CH];
  uint32_t cls_min, cls_max;
  uint32_t hkey[TM_HASH];
  uint16_t hhead[TM_HASH];
  uint16_t hnext[JT_CH];
  unsigned long long cnt[5];
  uint32_t seg_lo, seg_hi;
};

template <int F, bool T32>
__global__ void __launch_bounds__(JT_THREADS, 3)
k_join_ring(const __grid_constant__ JoinParams p, uint32_t *__restrict__ big_list, unsigned int *__restrict__ big_count,
            unsigned long long *__restrict__ seg_cursor) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  RingSmem &sm = *reinterpret_cast<RingSmem *>(smem_raw);
  const unsigned tid = threadIdx.x, lane = tid & 31;
  if (tid < 5) sm.cnt[tid] = 0;
  TileTotals w;
  const uint64_t W = (uint64_t)p.window_ns;
  const uint64_t Wi = W < (uint64_t)(500 * kMs) ? W : (uint64_t)(500 * kMs);
  // confidence falls with the tier (1.0, 0.9, 0.8, 0.65), so "confidence >= threshold" is "tier <= cut"
  const int cut = (1.0 >= p.threshold) + (0.9 >= p.threshold) + (0.8 >= p.threshold) + (0.65 >= p.threshold);
  const unsigned long long n_sp = p.n_spans;

  for (;;) {
    __syncthreads();   // everyone has read the previous segment's bounds
    if (tid == 0) {
      // a share of what is left (half of an even split), at least one chunk: large segments first, small ones at the end
      const unsigned long long cur = *reinterpret_cast<volatile unsigned long long *>(seg_cursor);
      const unsigned long long rem = cur < n_sp ? n_sp - cur : 0;
      unsigned long long take = rem / (2ull * gridDim.x);
      take = take > p.seg_max ? p.seg_max : take;
      take = take < (unsigned long long)JT_CH ? (unsigned long long)JT_CH : take;
      const unsigned long long start = atomicAdd(seg_cursor, take);
      sm.seg_lo = (uint32_t)(start < n_sp ? start : n_sp);
      sm.seg_hi = (uint32_t)(start + take < n_sp ? start + take : n_sp);
    }
    __syncthreads();
    size_t q = sm.seg_lo;
    const size_t q_end = sm.seg_hi;
    if (q >= q_end) break;
    uint32_t res_hi = 0;   // sorted positions below res_hi (and at or above the last tile's start) are resident
    bool have = false;

    while (q < q_end) {
      const int nq = (int)min((size_t)JT_CH, q_end - q);
      // ---- stage the next spans of the segment: input index, ranges, record; clear the hash and the per-span counts
      if ((int)tid < nq) {
        sm.rng[tid] = p.ranges[q + tid];
        sm.sidx[tid] = p.span_perm ? p.span_perm[q + tid] : (uint32_t)(q + tid);
      }
      // the whole-quanta / +-250 ms positions of the span this thread serves stay in its registers (2 KB of shared memory
      // that the note lists need)
      const uint4 r2 = (int)(tid >> 1) < nq ? __ldg(p.ranges2 + q + (tid >> 1)) : make_uint4(0u, 0u, 0u, 0u);
      if ((int)tid < nq) {
        const U8 u = ld_nc_v8(p.spans + sm.sidx[tid]);
        sm.span[tid][0] = u.a;
        sm.span[tid][1] = u.b;
      }
      sm.hkey[tid] = 0;            // JT_THREADS == TM_HASH
      sm.hhead[tid] = TM_END;
      if ((int)tid < JT_CH) { sm.n_note[tid] = 0; sm.n_tr[tid] = 0; sm.n_tr_in[tid] = 0; }
      if (tid == 0) { sm.cls_min = 0xffffffffu; sm.cls_max = 0u; }
      __syncthreads();
      // ---- hash of the staged spans' non-zero trace ids; spans that share an id are chained
      if ((int)tid < nq) {
        const uint32_t tr = sm.span[tid][1].y;
        if (tr != 0) {
          uint32_t slot = trace_slot(tr);
          for (;;) {
            const uint32_t prev = atomicCAS(&sm.hkey[slot], 0u, tr);
            if (prev == 0u || prev == tr) break;
            slot = (slot + 1) & (TM_HASH - 1);
          }
          unsigned short *hp = &sm.hhead[slot];
          unsigned short old = *hp, assumed;
          do {
            assumed = old;
            sm.hnext[tid] = assumed;
            old = atomicCAS(hp, assumed, (unsigned short)tid);
          } while (old != assumed);
        }
      }
      // ---- the tile: the staged spans whose outer range ends within RG_CAP positions of the first one's start
      const uint32_t A = sm.rng[0].x;
      const bool fits = (int)tid < nq && sm.rng[tid].y - A <= (uint32_t)RG_CAP;   // ranges are monotone: a prefix fits
      const int n_fit = __syncthreads_count(fits);
      if (n_fit == 0) {
        // this span's own range is larger than the ring: leave it to k_join_big (the ring keeps what it holds)
        if (tid == 0) big_list[atomicAdd(big_count, 1u)] = (uint32_t)q;
        q += 1;
        continue;
      }
      const uint32_t Bend = sm.rng[n_fit - 1].y;
      // positions [A, keep_hi) are resident already (ranges never move backwards inside a segment)
      const uint32_t keep_hi = have && res_hi > A ? min(res_hi, Bend) : A;
      uint32_t cmin = 0xffffffffu, cmax = 0u;
      // tier 1 for one tile event: probe the hash of the staged spans' trace ids, note the spans it falls into
      auto probe_trace = [&](const uint32_t g, const uint32_t tr, const int64_t e_ts) {
        if (tr == 0) return;
        uint32_t slot = trace_slot(tr);
        uint32_t key;
        while ((key = sm.hkey[slot]) != 0u && key != tr) slot = (slot + 1) & (TM_HASH - 1);
        if (key != tr) return;
        for (uint32_t k = sm.hhead[slot]; k != TM_END; k = sm.hnext[k]) {
          if ((int)k >= n_fit) continue;                              // that span belongs to the next tile
          const uint4 rg = sm.rng[k];
          if (g < rg.x || g >= rg.y) continue;                        // outside its range: cannot be within the window
          const int64_t s_ts = (int64_t)(((uint64_t)sm.span[k][0].y << 32) | sm.span[k][0].x);
          const uint64_t adt = lsg_absdiff(s_ts, e_ts);
          if (adt > W) continue;                                      // outer window, dns.go:54-56
          atomicAdd(&sm.n_tr[k], 1u);
          if (adt <= Wi) atomicAdd(&sm.n_tr_in[k], 1u);
          if (cut >= 1) {
            const uint32_t at = atomicAdd(&sm.n_note[k], 1u);
            if (at < RG_LIST) sm.note[at][k] = (uint16_t)((LSG_TIER_TRACE << 12) | (g & RG_MASK));
          }
        }
      };
      {
        // ---- gather what is not resident: first the thread's permutation entries (coalesced), then the records as
        //      independent 256-bit loads, LSG_JT_DEPTH at a time
        uint32_t src[RG_PER];
#pragma unroll
        for (int u = 0; u < RG_PER; ++u) {
          const uint32_t g = keep_hi + tid + u * JT_THREADS;
          src[u] = g < Bend ? __ldg(p.ev_perm + g) : 0xffffffffu;
        }
#pragma unroll
        for (int ub = 0; ub < RG_PER; ub += LSG_JT_DEPTH) {
          U8 r[LSG_JT_DEPTH];
#pragma unroll
          for (int u = 0; u < LSG_JT_DEPTH; ++u)
            if (src[ub + u] != 0xffffffffu) r[u] = ld_nc_v8(p.events_in + src[ub + u]);
#pragma unroll
          for (int u = 0; u < LSG_JT_DEPTH; ++u) {
            if (src[ub + u] == 0xffffffffu) continue;
            const uint32_t g = keep_hi + tid + (ub + u) * JT_THREADS;
            const uint32_t sl = g & RG_MASK;
            const uint32_t row = r[u].b.x;
            const uint2 pc = row < p.n_idents ? __ldg(p.ident_pc + row) : make_uint2(0u, 0u);
            const int64_t e_ts = (int64_t)(((uint64_t)r[u].a.y << 32) | r[u].a.x);
            sm.ts[sl] = e_ts;
            sm.pod[sl] = pc.x;
            sm.cls[sl] = pc.y;
            sm.pid[sl] = r[u].b.y;
            sm.conn[sl] = r[u].b.z;
            sm.trsig[sl] = r[u].b.w;
            cmin = min(cmin, pc.y);
            cmax = max(cmax, pc.y);
            probe_trace(g, r[u].b.w >> 4, e_ts);   // tier 1 while the record is still in registers
          }
        }
        // ---- what stayed from the previous tile meets this tile's spans: class range and tier 1 from shared memory
        for (uint32_t g = A + tid; g < keep_hi; g += JT_THREADS) {
          const uint32_t sl = g & RG_MASK;
          const uint32_t c = sm.cls[sl];
          cmin = min(cmin, c);
          cmax = max(cmax, c);
          probe_trace(g, sm.trsig[sl] >> 4, sm.ts[sl]);
        }
      }
      cmin = __reduce_min_sync(0xffffffffu, cmin);
      cmax = __reduce_max_sync(0xffffffffu, cmax);
      if (lane == 0 && cmin <= cmax) { atomicMin(&sm.cls_min, cmin); atomicMax(&sm.cls_max, cmax); }
      __syncthreads();
      // ---- tiers 2-4: thread (2k + h) walks half h of span k's inner range
      const bool uniform = sm.cls_min == sm.cls_max;
      const int k = (int)(tid >> 1);
      const unsigned h = tid & 1;
      const bool live = k < n_fit;
      unsigned n_match = 0, n_low = 0, n_unlisted = 0;
      SpanRec s;
      s.ts = 0; s.ident = 0; s.pid = 0; s.conn = 0; s.trace = 0; s.incident = LSG_NONE;
      if (live) {
        s = span_from_halves(sm.span[k][0], sm.span[k][1]);
        const uint4 rg = sm.rng[k];
        // `fanout` trace-id matches (confidence 1.0) already noted: a tier 2-4 candidate can be counted but never kept
        const bool full = cut >= 1 && sm.n_tr[k] >= (uint32_t)p.fanout;
        if (rg.z < rg.w && rg.z != 0xffffffffu) {
          const uint2 spc = __ldg(p.ident_pc + s.ident);
          // an empty pod / service / node never matches (dns.go:61-73); classes are <= n_idents, so ~0 is no event's class
          const bool pod_ok = spc.x != 0;
          const uint32_t s_pod = spc.x, s_cls = spc.y != 0 ? spc.y : 0xffffffffu;
          const bool pid_ok = s.pid != 0, conn_ok = s.conn != 0;
          PairSpan ps;
          ps.ts = s.ts; ps.s_lo = (uint32_t)(uint64_t)s.ts; ps.pod = spc.x; ps.cls = s_cls; ps.pid = s.pid; ps.conn = s.conn;
          ps.no_trace = s.trace != 0 ? s.trace : 0xffffffffu;   // event trace ids are < 2^28
          ps.pod_ok = pod_ok; ps.pid_ok = pid_ok; ps.conn_ok = conn_ok; ps.full = full;
          const TileCols cols{sm.ts, sm.pod, sm.cls, sm.pid, sm.conn, sm.trsig};
          const auto ring = [](uint32_t g) { return g & RG_MASK; };
          if (cut < 4 && uniform) {
            // one (service, node) class in the whole tile and a service+node match is below the threshold: tier 4 is
            // counted (events inside the inner window - trace-id matches among them - tier 2/3 matches), see k_join_tiles
            uint32_t n_in = h ? 0u : r2.y - r2.x;
            const uint32_t eb = h ? r2.y : rg.z, ee = h ? rg.w : r2.x;      // lane 0: left edge, lane 1: right edge
            const uint32_t *ts_lo = reinterpret_cast<const uint32_t *>(sm.ts);
            for (uint32_t g = eb; g < ee; ++g) {
              const uint32_t sl = g & RG_MASK;
              if (T32) {
                const int32_t d = (int32_t)(ts_lo[2 * sl] - ps.s_lo);
                n_in += (uint32_t)(d < 0 ? -d : d) <= (uint32_t)Wi;
              } else {
                n_in += lsg_absdiff(s.ts, sm.ts[sl]) <= Wi;
              }
            }
            const uint32_t m2 = r2.z + ((r2.w - r2.z) >> 1);
            const unsigned before = n_match;
            pair_scan<T32, false, RG_LIST>(cols, h ? m2 : r2.z, h ? r2.w : m2, ring, ps, Wi, cut, &sm.n_note[k], &sm.note[0][k], n_match,
                                           n_low, n_unlisted);
            if (s_cls == sm.cls_min) {   // the span belongs to the tile's class: what is left inside the window is tier 4
              uint32_t n4 = n_in - (n_match - before);                             // may wrap per lane; the two-lane sum does not
              if (!h) n4 -= sm.n_tr_in[k];
              n_match += n4;
              n_low += n4;
            }
          } else {
            const uint32_t mid = rg.z + ((rg.w - rg.z) >> 1);
            pair_scan<T32, true, RG_LIST>(cols, h ? mid : rg.z, h ? rg.w : mid, ring, ps, Wi, cut, &sm.n_note[k], &sm.note[0][k], n_match,
                                          n_low, n_unlisted);
          }
        }
      }
      __syncwarp();      // the two lanes of a span sit in one warp: their notes are visible to each other now
      // ---- order the span's notes, write its outputs (span_epilogue, join.cu)
      {
        const TileCols cols{sm.ts, sm.pod, sm.cls, sm.pid, sm.conn, sm.trsig};
        span_epilogue<F, true, RG_LIST>(p, cols, &sm.note[0][live ? k : 0], live ? sm.n_note[k] : 0u, live ? sm.n_tr[k] : 0u, live,
                                        live && sm.rng[k].z == 0xffffffffu, s, live ? (size_t)sm.sidx[k] : (size_t)0, (uint32_t)(q + k), h,
                                        lane, cut, n_match, n_low, n_unlisted,
                                        [A](uint32_t sl) { return A + ((sl - A) & RG_MASK); },   // the resident position in that slot
                                        [](uint32_t g) { return g & RG_MASK; }, big_list, big_count, w);
      }
      __syncthreads();   // every thread is done with the staged spans and with the slots the next tile overwrites
      q += n_fit;
      res_hi = Bend;
      have = true;
    }
  }
  flush_totals(w, sm.cnt, p.counters);
}
