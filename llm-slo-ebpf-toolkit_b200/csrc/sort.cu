// sort.cu — K-sort: events (and spans) grouped by (join group, quantised time).
//
// What the reference does here: nothing — ProcessBatch loops every span over every signal
// (ebpfcorrelator/processor.go:29-49, correlator.go:72-90).  Grouping is how this library reaches the
// same match sets in O(E + S*w) instead of O(S*E).
//
// Keys are (grp << tbits) | ((ts - ts_min) >> shift).  The join never needs a total order in time (it
// selects candidates by an explicit (tier, |dt|, input index) order), only that a span's window is one
// contiguous range, so the time quantum 2^shift is as large as the window allows (<= window/32) and the
// key shrinks from ~50 bits to 23-30.  `exact` (lsg_sort_events_dev) uses shift 0.
//
// The time axis of the key comes from the spans (first live span - window .. last live span + window: an event
// outside it can match nothing), so nothing reads the events before the sort's own first pass.
//
// Two ways to group, both ending in a permutation (sorted position -> input index).  The pipeline stops there: the
// join gathers its tiles through the permutation.  A gather that moves the 32-byte records (random reads, sequential
// writes; random 32-byte writes measured 2.6x slower) only runs for lsg_sort_events_dev and with LSG_JOIN_SORTED_COPY:
//   counting sort  (key space <= 2^25, not exact)  one counter per key in an L2-resident table:
//                  k_count_keys (key + rank by returning atomic) -> exclusive scan -> k_place_events
//                  (permutation).  Order inside one key = whatever the atomics gave.
//   LSD radix sort (larger key spaces, spans, the exact entry point)  stable; digits of 8-10 bits that share the
//                  key's bits evenly (29 bits: three passes); per pass: per-tile digit histogram (shared-memory
//                  atomics) -> per-digit scan over tiles -> scatter of (key, index) pairs.  Ranking inside
//                  a tile is warp-synchronous (peer masks from one ballot per digit bit, one 16-bit shared-memory
//                  counter row per warp, no atomics); items are staged in shared memory in digit order so that the
//                  global writes are coalesced runs.  No inter-CTA spin-waits: tiles only communicate
//                  through the scanned histogram.
#include "common.cuh"

namespace {

constexpr int RADIX_MAX_BITS = 10;   // widest digit of a radix pass
constexpr int SORT_BLOCK = 512;
constexpr int SORT_IPT = 16;
constexpr int SORT_TILE = SORT_BLOCK * SORT_IPT;  // 8192 keys per CTA
constexpr int SORT_NW = SORT_BLOCK / 32;

// ---- statistics that fix the key layout ----------------------------------------------------------
using EventStats = LsgBatchStats;

__global__ void k_ident_stats(const lsg_ident *__restrict__ idents, uint32_t n, EventStats *st) {
  unsigned int m = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) m = max(m, idents[i].grp);
  for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m) atomicMax(&st->max_grp, m);
}

// time range of the spans that can match anything; a span whose ident is outside the table is a caller error (bit 1)
__global__ void __launch_bounds__(256)
k_span_stats(const lsg_span *__restrict__ spans, size_t n, const lsg_ident *__restrict__ idents, uint32_t n_idents,
             EventStats *st) {
  long long tmin = INT64_MAX, tmax = INT64_MIN;
  bool bad = false;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const SpanRec s = load_span(spans + i);
    if (s.ident >= n_idents) { bad = true; continue; }
    if (s.ts == LSG_TS_ZERO || load_ident(idents, s.ident).grp == 0) continue;
    tmin = min(tmin, (long long)s.ts);
    tmax = max(tmax, (long long)s.ts);
  }
  // warp -> block -> one pair of atomics per block (every warp hitting the same two words serialises in L2)
  __shared__ long long s_min[8], s_max[8];
  __shared__ int s_bad;
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  for (int o = 16; o; o >>= 1) {
    tmin = min(tmin, __shfl_xor_sync(0xffffffffu, tmin, o));
    tmax = max(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
  }
  const bool warp_bad = __any_sync(0xffffffffu, bad);
  if ((threadIdx.x & 31) == 0) {
    s_min[threadIdx.x >> 5] = tmin;
    s_max[threadIdx.x >> 5] = tmax;
    if (warp_bad) s_bad = 1;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { tmin = min(tmin, s_min[w]); tmax = max(tmax, s_max[w]); }
    if (s_bad) atomicOr(&st->bad, 2u);
    if (tmin <= tmax) {
      atomicMin(&st->ts_min, tmin);
      atomicMax(&st->ts_max, tmax);
    }
  }
}

__device__ __forceinline__ bool event_live(const EventRec &e, const lsg_ident *idents, uint32_t n_idents, uint32_t *grp,
                                           bool *bad) {
  *grp = 0;
  if (e.ident >= n_idents) { *bad = true; return false; }
  if ((e.trace_sig & 15u) >= LSG_NSIG) return false;  // signalAttrKey: unsupported (correlator.go:73-77)
  if (e.ts == LSG_TS_ZERO) return false;               // zero Time never matches (dns.go:108)
  *grp = load_ident(idents, e.ident).grp;
  return *grp != 0;
}

// the caller stated the spans' time range and the largest group (LsgKeyHint): nobody waits for the statistics, they are
// only checked against the statement; a violated one voids the batch (bit 2 of `bad`, reported with the counters)
__global__ void k_check_hint(EventStats *st, long long lo, long long hi, unsigned int max_grp) {
  if (st->ts_min <= st->ts_max && (st->ts_min < lo || st->ts_max > hi)) st->bad |= 4u;
  if (st->max_grp > max_grp) st->bad |= 4u;
}

// the sort's own look at an event: its key (or "dead"), and what the batch statistics need from it
struct EventTally {
  unsigned int unsup = 0;
  bool bad = false;
  __device__ __forceinline__ void flush(EventStats *st) {
    if (!st) return;
    unsup = __reduce_add_sync(0xffffffffu, unsup);
    const bool any_bad = __any_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0) {
      if (unsup) atomicAdd(&st->n_unsupported, (unsigned long long)unsup);
      if (any_bad) atomicOr(&st->bad, 1u);
    }
  }
};
__device__ __forceinline__ bool event_key(const EventRec &e, const lsg_ident *idents, uint32_t n_idents, const KeySpec &ks,
                                          EventTally &t, uint64_t *key) {
  if ((e.trace_sig & 15u) >= LSG_NSIG) t.unsup++;
  uint32_t g;
  if (!event_live(e, idents, n_idents, &g, &t.bad)) return false;
  if (e.ts < ks.ts_min || e.ts > ks.ts_hi) return false;   // further than the window from every span
  *key = lsg_make_key(ks, g, e.ts);
  return *key < ks.dead;   // a group beyond the stated maximum (key hint that does not hold): never index past the tables
}

__global__ void __launch_bounds__(256)
k_event_stats(const lsg_event *__restrict__ events, size_t n, const lsg_ident *__restrict__ idents, uint32_t n_idents,
              EventStats *st) {
  long long tmin = INT64_MAX, tmax = INT64_MIN;
  unsigned int live = 0, unsup = 0;
  bool bad = false;
  auto look = [&](const EventRec &e) {
    uint32_t g;
    if ((e.trace_sig & 15u) >= LSG_NSIG) unsup++;
    if (event_live(e, idents, n_idents, &g, &bad)) {
      live++;
      tmin = min(tmin, (long long)e.ts);
      tmax = max(tmax, (long long)e.ts);
    }
  };
  // streaming read of every record: four 256-bit loads in flight per thread
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    EventRec e[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) e[u] = load_event(events + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) look(e[u]);
  }
  for (; i < n; i += stride) look(load_event(events + i));
  for (int o = 16; o; o >>= 1) {
    tmin = min(tmin, __shfl_xor_sync(0xffffffffu, tmin, o));
    tmax = max(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
    live += __shfl_xor_sync(0xffffffffu, live, o);
    unsup += __shfl_xor_sync(0xffffffffu, unsup, o);
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(&st->bad, 1u);
  if ((threadIdx.x & 31) == 0) {
    if (live) {
      atomicMin(&st->ts_min, tmin);
      atomicMax(&st->ts_max, tmax);
      atomicAdd(&st->n_live, (unsigned long long)live);
    }
    if (unsup) atomicAdd(&st->n_unsupported, (unsigned long long)unsup);
  }
}

template <typename KeyT>
__global__ void __launch_bounds__(256)
k_event_keys(const lsg_event *__restrict__ events, size_t n, const lsg_ident *__restrict__ idents, uint32_t n_idents,
             KeySpec ks, KeyT *__restrict__ keys, EventStats *st) {
  EventTally tally;
#pragma unroll 4
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    EventRec e = load_event(events + i);
    uint64_t k = ks.dead;
    event_key(e, idents, n_idents, ks, tally, &k);
    keys[i] = (KeyT)k;
  }
  tally.flush(st);
}

template <typename KeyT>
__global__ void __launch_bounds__(256)
k_span_keys(const lsg_span *__restrict__ spans, size_t n, const lsg_ident *__restrict__ idents, uint32_t n_idents, KeySpec ks,
            int64_t ts_max, KeyT *__restrict__ keys) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    SpanRec s = load_span(spans + i);
    KeyT k = (KeyT)ks.dead;
    const uint32_t g = s.ident < n_idents ? load_ident(idents, s.ident).grp : 0u;   // a bad ident is reported elsewhere
    if (g != 0 && s.ts != LSG_TS_ZERO) {
      int64_t t = s.ts < ks.ts_min ? ks.ts_min : (s.ts > ts_max ? ts_max : s.ts);   // processing order only
      k = (KeyT)lsg_make_key(ks, g, t);
    }
    keys[i] = k;
  }
}

// ---- one radix pass ------------------------------------------------------------------------------
// The digit is BITS wide (8, 9 or 10): the passes share the key's bits evenly, so 17-20 bits take two passes and
// 25-30 bits three (with 8-bit digits alone 29 bits would take four, each a read and a write of every pair).
// lanes of the warp whose digit equals this lane's: BITS ballots instead of match.any, which is issued through the
// (slow, shared) MIO queue on sm_100 and throttled both radix kernels
template <int BITS>
__device__ __forceinline__ unsigned peers_of_digit(const unsigned d) {
  unsigned peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < BITS; ++b) {
    const bool bit = (d >> b) & 1u;
    const unsigned bal = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? bal : ~bal;
  }
  return peers;
}

// Per-tile digit counts.  Shared-memory atomics on four copies of the counter row (a warp whose 32 keys all carry one
// digit adds 32 at once: the top digit of a batch with few groups); the ballot ranking of the scatter is not needed to
// count, and cost 3.6 ms per pass over 10^9 keys against 0.7 ms for the read.
template <typename KeyT, int BITS>
__global__ void __launch_bounds__(SORT_BLOCK)
k_radix_hist(const KeyT *__restrict__ keys, size_t n, int shift, uint32_t *__restrict__ tile_hist, uint32_t n_tiles) {
  constexpr int R = 1 << BITS;
  __shared__ uint32_t s_hist[4][R];
  for (int i = threadIdx.x; i < 4 * R; i += SORT_BLOCK) (&s_hist[0][0])[i] = 0;
  __syncthreads();
  const size_t base = (size_t)blockIdx.x * SORT_TILE;
  const unsigned lane = threadIdx.x & 31;
  uint32_t *mine = s_hist[(threadIdx.x >> 5) & 3];
  KeyT k[SORT_IPT];
#pragma unroll
  for (int i = 0; i < SORT_IPT; ++i) {
    const size_t pos = base + (size_t)i * SORT_BLOCK + threadIdx.x;
    k[i] = pos < n ? keys[pos] : (KeyT)0;
  }
#pragma unroll
  for (int i = 0; i < SORT_IPT; ++i) {
    const size_t pos = base + (size_t)i * SORT_BLOCK + threadIdx.x;
    const bool valid = pos < n;
    const unsigned d = (unsigned)((k[i] >> shift) & (R - 1));
    const unsigned d0 = __shfl_sync(0xffffffffu, d, 0);
    if (__all_sync(0xffffffffu, valid && d == d0)) {
      if (lane == 0) atomicAdd(&mine[d0], 32u);
    } else if (valid) {
      atomicAdd(&mine[d], 1u);
    }
  }
  __syncthreads();
  for (int d = threadIdx.x; d < R; d += SORT_BLOCK)
    tile_hist[(size_t)d * n_tiles + blockIdx.x] = s_hist[0][d] + s_hist[1][d] + s_hist[2][d] + s_hist[3][d];
}

// exclusive scan of each digit's row over the tiles (in place); one CTA per digit
__global__ void __launch_bounds__(1024)
k_radix_scan_tiles(uint32_t *__restrict__ tile_hist, uint32_t n_tiles, uint32_t *__restrict__ digit_total) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  uint32_t *row = tile_hist + (size_t)blockIdx.x * n_tiles;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t base = 0; base < n_tiles; base += 1024) {
    uint32_t i = base + threadIdx.x;
    uint32_t v = i < n_tiles ? row[i] : 0, x = v;
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = s_warp[lane], z = w;
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, z, o);
        if (lane >= (unsigned)o) z += y;
      }
      s_warp[lane] = z - w;  // exclusive over warps
    }
    __syncthreads();
    uint32_t carry = s_carry;
    if (i < n_tiles) row[i] = carry + s_warp[warp] + x - v;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = carry + s_warp[warp] + x;
    __syncthreads();
  }
  if (threadIdx.x == 0) digit_total[blockIdx.x] = s_carry;
}

// exclusive scan of the digit totals: one CTA of as many threads as there are digits (256 .. 1024)
__global__ void __launch_bounds__(1024) k_radix_scan_digits(const uint32_t *__restrict__ digit_total, uint32_t *__restrict__ digit_base) {
  __shared__ uint32_t s_warp[32];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t v = digit_total[threadIdx.x], x = v;
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (unsigned)o) x += y;
  }
  if (lane == 31) s_warp[warp] = x;
  __syncthreads();
  uint32_t pre = 0;
  for (unsigned w = 0; w < warp; ++w) pre += s_warp[w];
  digit_base[threadIdx.x] = pre + x - v;
}

// shared memory of the scatter: per-warp digit counters (16 bits: a warp ranks 512 items, a tile holds 8192), the
// digits' tile-local and global bases, and the tile's pairs in digit order
template <typename KeyT, int BITS>
constexpr size_t scatter_smem_bytes() {
  return (size_t)SORT_NW * (1 << BITS) * 2 + (size_t)(1 << BITS) * 8 + 64 * 4 + (size_t)SORT_TILE * 4 + (size_t)SORT_TILE * sizeof(KeyT);
}

template <typename KeyT, int BITS>
__global__ void __launch_bounds__(SORT_BLOCK, 2)
k_radix_scatter(const KeyT *__restrict__ keys_in, const uint32_t *__restrict__ idx_in, KeyT *__restrict__ keys_out,
                uint32_t *__restrict__ idx_out, size_t n, int shift, const uint32_t *__restrict__ tile_off,
                const uint32_t *__restrict__ digit_base, uint32_t n_tiles) {
  constexpr int R = 1 << BITS;
  constexpr int DPT = R > SORT_BLOCK ? R / SORT_BLOCK : 1;   // digits per thread in the tile-level scan
  extern __shared__ __align__(16) unsigned char smem[];
  uint32_t *bin_off = reinterpret_cast<uint32_t *>(smem);   // [R] first tile-local slot of each digit
  uint32_t *gbase = bin_off + R;                            // [R] global position of tile-local slot 0, per digit
  uint32_t *s_scan = gbase + R;                             // [64]
  uint32_t *s_idx = s_scan + 64;                            // [SORT_TILE]
  KeyT *s_keys = reinterpret_cast<KeyT *>(s_idx + SORT_TILE);  // [SORT_TILE]
  uint16_t *warp_cnt = reinterpret_cast<uint16_t *>(s_keys + SORT_TILE);  // [SORT_NW][R]

  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const size_t base = (size_t)blockIdx.x * SORT_TILE;
  const uint32_t count = (uint32_t)min((size_t)SORT_TILE, n - base);
  for (int i = tid; i < SORT_NW * R / 2; i += SORT_BLOCK) reinterpret_cast<uint32_t *>(warp_cnt)[i] = 0;
  __syncthreads();

  // warp-striped load: warp w owns tile slots [w*32*IPT, (w+1)*32*IPT), item i of lane l is slot i*32+l
  KeyT k[SORT_IPT];
  uint16_t rank[SORT_IPT];
  const size_t wbase = base + (size_t)warp * 32 * SORT_IPT;
#pragma unroll
  for (int i = 0; i < SORT_IPT; ++i) {
    size_t pos = wbase + (size_t)i * 32 + lane;
    k[i] = pos < n ? keys_in[pos] : (KeyT)~(KeyT)0;   // padding sorts behind every real key of the tile (stable)
  }
  // stable rank of each item among the items of its warp with the same digit
  uint16_t *my_cnt = warp_cnt + warp * R;
  const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < SORT_IPT; ++i) {
    unsigned d = (unsigned)((k[i] >> shift) & (R - 1));
    unsigned m = peers_of_digit<BITS>(d);
    int leader = __ffs(m) - 1;
    uint32_t old = 0;
    if ((int)lane == leader) {
      old = my_cnt[d];
      my_cnt[d] = (uint16_t)(old + __popc(m));
    }
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[i] = (uint16_t)(old + __popc(m & lt_mask));
    __syncwarp();
  }
  __syncthreads();
  // per digit: exclusive scan over the warps, then over the digits (thread t owns digits [t*DPT, (t+1)*DPT))
  uint32_t tot[DPT];
  uint32_t mine = 0;
#pragma unroll
  for (int q = 0; q < DPT; ++q) {
    const unsigned d = tid * DPT + q;
    tot[q] = 0;
    if (d < (unsigned)R) {
      uint32_t run = 0;
#pragma unroll
      for (int w = 0; w < SORT_NW; ++w) {
        const uint32_t c = warp_cnt[w * R + d];
        warp_cnt[w * R + d] = (uint16_t)run;
        run += c;
      }
      tot[q] = run;
    }
    mine += tot[q];
  }
  uint32_t x = mine;
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (unsigned)o) x += y;
  }
  if (lane == 31) s_scan[warp] = x;
  __syncthreads();
  {
    uint32_t pre = 0;
    for (unsigned w = 0; w < warp; ++w) pre += s_scan[w];
    uint32_t excl = pre + x - mine;
#pragma unroll
    for (int q = 0; q < DPT; ++q) {
      const unsigned d = tid * DPT + q;
      if (d < (unsigned)R) {
        bin_off[d] = excl;
        gbase[d] = digit_base[d] + tile_off[(size_t)d * n_tiles + blockIdx.x] - excl;
        excl += tot[q];
      }
    }
  }
  __syncthreads();
  // the payload (input index) is read only now, so that it never sits in registers during the ranking
#pragma unroll
  for (int i = 0; i < SORT_IPT; ++i) {
    const size_t pos = wbase + (size_t)i * 32 + lane;
    const uint32_t v = pos < n ? (idx_in ? idx_in[pos] : (uint32_t)pos) : 0u;
    unsigned d = (unsigned)((k[i] >> shift) & (R - 1));
    uint32_t lp = bin_off[d] + my_cnt[d] + rank[i];
    s_keys[lp] = k[i];
    s_idx[lp] = v;
  }
  __syncthreads();
  // tile is now in digit order in shared memory: runs of equal digits go to consecutive global slots
  for (uint32_t j = tid; j < count; j += SORT_BLOCK) {
    KeyT key = s_keys[j];
    unsigned d = (unsigned)((key >> shift) & (R - 1));
    size_t out = (size_t)gbase[d] + j;
    keys_out[out] = key;
    idx_out[out] = s_idx[j];
  }
}

__global__ void __launch_bounds__(256)
k_gather_events(const lsg_event *__restrict__ events, const uint32_t *__restrict__ perm, size_t n,
                const uint32_t *__restrict__ n_placed, lsg_event *__restrict__ sorted) {
  // one 256-bit load and one 256-bit store per record; eight records in flight per thread
  if (n_placed) n = min(n, (size_t)*n_placed);   // counting sort: only the live events have a permutation entry
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; j + 7 * stride < n; j += 8 * stride) {
    U8 r[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) r[u] = ld_nc_v8(events + perm[j + u * stride]);
#pragma unroll
    for (int u = 0; u < 8; ++u) st_v8(sorted + j + u * stride, r[u]);
  }
  for (; j < n; j += stride) st_v8(sorted + j, ld_nc_v8(events + perm[j]));
}

// the same gather into columns: what the join stages into shared memory (ts, ident, pid, conn, trace|signal) plus values
__global__ void __launch_bounds__(256)
k_gather_events_soa(const lsg_event *__restrict__ events, const uint32_t *__restrict__ perm, size_t n,
                    const uint32_t *__restrict__ n_placed, int64_t *__restrict__ ts,
                    double *__restrict__ val, uint32_t *__restrict__ ident, uint32_t *__restrict__ pid,
                    uint32_t *__restrict__ conn, uint32_t *__restrict__ trsig) {
  if (n_placed) n = min(n, (size_t)*n_placed);   // counting sort: only the live events have a permutation entry
  auto put = [&](const size_t j, const U8 r) {
    ts[j] = (int64_t)(((uint64_t)r.a.y << 32) | r.a.x);
    val[j] = __longlong_as_double((long long)(((uint64_t)r.a.w << 32) | r.a.z));
    ident[j] = r.b.x;
    pid[j] = r.b.y;
    conn[j] = r.b.z;
    trsig[j] = r.b.w;
  };
  // one 256-bit load per record, eight records in flight per thread; the column stores are coalesced
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; j + 7 * stride < n; j += 8 * stride) {
    U8 r[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) r[u] = ld_nc_v8(events + perm[j + u * stride]);
#pragma unroll
    for (int u = 0; u < 8; ++u) put(j + u * stride, r[u]);
  }
  for (; j < n; j += stride) put(j, ld_nc_v8(events + perm[j]));
}

// ---- counting sort (small key spaces) --------------------------------------------------------------------------
// With the coarse time quantum the whole key space of a batch is a few million values: one counter per key fits
// in L2, so "sort" is: count the keys (L2 atomics), exclusive scan, place every record at a cursor fetch-add.
// Three light passes instead of keys + 3 x (histogram, scan, scatter), and a span's event range becomes
// two table reads.  The order of the events inside one key is whatever the atomics give: the join does not care
// (it selects by an explicit total order), and lsg_sort_events_dev never takes this path.
constexpr uint32_t COUNTING_MAX_KEYS = 1u << 25;   // 128 MB of counters at most

// pass 1: one look at every record.  The returning atomic gives the event's rank inside its key's run; key and
// rank are kept (8 B/event) so that the placement below neither repeats the atomic nor reads the records again.
__global__ void __launch_bounds__(256)
k_count_keys(const lsg_event *__restrict__ events, size_t n, const lsg_ident *__restrict__ idents, uint32_t n_idents,
             KeySpec ks, uint32_t *__restrict__ bins, uint2 *__restrict__ key_rank, EventStats *st) {
  // the returning atomic has L2 round-trip latency: four events per thread in flight
  EventTally tally;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    EventRec e[4];
    uint2 kr[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) e[u] = load_event(events + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      uint64_t k;
      kr[u] = make_uint2(0xffffffffu, 0u);   // not placed: matches nothing
      if (event_key(e[u], idents, n_idents, ks, tally, &k)) kr[u].x = (uint32_t)k;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (kr[u].x != 0xffffffffu) kr[u].y = atomicAdd(&bins[kr[u].x], 1u);
#pragma unroll
    for (int u = 0; u < 4; ++u) key_rank[i + u * stride] = kr[u];
  }
  for (; i < n; i += stride) {
    const EventRec e = load_event(events + i);
    uint64_t k;
    uint2 kr = make_uint2(0xffffffffu, 0u);
    if (event_key(e, idents, n_idents, ks, tally, &k)) {
      kr.x = (uint32_t)k;
      kr.y = atomicAdd(&bins[kr.x], 1u);
    }
    key_rank[i] = kr;
  }
  tally.flush(st);
}

// pass 3 (after the scan turned the counts into run starts): position = start of the key's run + rank.
// Only the permutation is scattered (4-byte writes into an array that stays in L2); the 32-byte records are moved
// by whoever needs them (the join's tile loader, or the gather): random READS of records are cheap, random 32-byte
// WRITES into a 320 MB output are not (measured: 0.74 ms for 10 M records against 0.25 ms for a gather).
__global__ void __launch_bounds__(256)
k_place_events(const uint2 *__restrict__ key_rank, size_t n, const uint32_t *__restrict__ bins, uint32_t *__restrict__ perm) {
  // two dependent random accesses per event (run start, then the permutation slot): four events per thread in flight
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    uint2 kr[4];
    uint32_t start[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) kr[u] = key_rank[i + u * stride];
#pragma unroll
    for (int u = 0; u < 4; ++u) start[u] = kr[u].x != 0xffffffffu ? __ldg(bins + kr[u].x) : 0u;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (kr[u].x != 0xffffffffu) perm[start[u] + kr[u].y] = (uint32_t)(i + u * stride);
  }
  for (; i < n; i += stride) {
    const uint2 kr = key_rank[i];
    if (kr.x != 0xffffffffu) perm[__ldg(bins + kr.x) + kr.y] = (uint32_t)i;
  }
}

// exclusive scan of n 32-bit counters in place: block sums, scan of the sums, local scan + offset
constexpr int SCAN_BLOCK = 256, SCAN_IPT = 16, SCAN_TILE = SCAN_BLOCK * SCAN_IPT;

__global__ void __launch_bounds__(SCAN_BLOCK)
k_scan_reduce(const uint32_t *__restrict__ data, size_t n, uint32_t *__restrict__ block_sums) {
  __shared__ uint32_t s_w[SCAN_BLOCK / 32];
  const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_IPT;
  uint32_t t = 0;
  if (base + SCAN_IPT <= n) {
#pragma unroll
    for (int q = 0; q < SCAN_IPT / 4; ++q) {
      const uint4 v = *reinterpret_cast<const uint4 *>(data + base + 4 * q);
      t += v.x + v.y + v.z + v.w;
    }
  } else {
    for (int q = 0; q < SCAN_IPT; ++q)
      if (base + q < n) t += data[base + q];
  }
  t = __reduce_add_sync(0xffffffffu, t);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t b = 0;
    for (int w = 0; w < SCAN_BLOCK / 32; ++w) b += s_w[w];
    block_sums[blockIdx.x] = b;
  }
}

__global__ void __launch_bounds__(1024) k_scan_sums(uint32_t *__restrict__ sums, uint32_t n) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t base = 0; base < n; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t v = i < n ? sums[i] : 0;
    uint32_t x = v;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
      const uint32_t w = s_warp[lane];
      uint32_t z = w;
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, z, o);
        if (lane >= (unsigned)o) z += y;
      }
      s_warp[lane] = z - w;
    }
    __syncthreads();
    const uint32_t carry = s_carry;
    if (i < n) sums[i] = carry + s_warp[warp] + x - v;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = carry + s_warp[warp] + x;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(SCAN_BLOCK)
k_scan_apply(uint32_t *__restrict__ data, size_t n, const uint32_t *__restrict__ block_sums) {
  __shared__ uint32_t s_w[SCAN_BLOCK / 32];
  const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_IPT;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t v[SCAN_IPT];
  const bool full = base + SCAN_IPT <= n;
  if (full) {
#pragma unroll
    for (int q = 0; q < SCAN_IPT / 4; ++q) {
      const uint4 u = *reinterpret_cast<const uint4 *>(data + base + 4 * q);
      v[4 * q] = u.x; v[4 * q + 1] = u.y; v[4 * q + 2] = u.z; v[4 * q + 3] = u.w;
    }
  } else {
#pragma unroll
    for (int q = 0; q < SCAN_IPT; ++q) v[q] = base + q < n ? data[base + q] : 0;
  }
  uint32_t t = 0;
#pragma unroll
  for (int q = 0; q < SCAN_IPT; ++q) { const uint32_t c = v[q]; v[q] = t; t += c; }   // exclusive within the thread
  uint32_t x = t;
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= (unsigned)o) x += y;
  }
  if (lane == 31) s_w[warp] = x;
  __syncthreads();
  uint32_t pre = block_sums[blockIdx.x] + x - t;
  for (unsigned w = 0; w < warp; ++w) pre += s_w[w];
  if (full) {
#pragma unroll
    for (int q = 0; q < SCAN_IPT / 4; ++q)
      *reinterpret_cast<uint4 *>(data + base + 4 * q) = make_uint4(pre + v[4 * q], pre + v[4 * q + 1], pre + v[4 * q + 2], pre + v[4 * q + 3]);
  } else {
#pragma unroll
    for (int q = 0; q < SCAN_IPT; ++q)
      if (base + q < n) data[base + q] = pre + v[q];
  }
}

inline unsigned grid_for(lsg_ctx *ctx, size_t n, int block, int per_sm) {
  size_t blocks = (n + block - 1) / block;
  size_t cap = (size_t)ctx->sm_count * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  return (unsigned)blocks;
}

// one pass: histogram, scans, scatter
template <typename KeyT, int BITS>
int radix_pass(lsg_ctx *ctx, const KeyT *kin, const uint32_t *iin, KeyT *kout, uint32_t *iout, size_t n, int shift, uint32_t n_tiles,
               uint32_t *tile_hist, uint32_t *digit_total, uint32_t *digit_base, int prof_base) {
  constexpr int R = 1 << BITS;
  constexpr size_t smem = scatter_smem_bytes<KeyT, BITS>();
  LSG_CUDA(ctx, cudaFuncSetAttribute(k_radix_scatter<KeyT, BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  {
    ProfScope ps(ctx, prof_base >= 0 ? prof_base : PK_HIST);
    k_radix_hist<KeyT, BITS><<<n_tiles, SORT_BLOCK, 0, ctx->stream>>>(kin, n, shift, tile_hist, n_tiles);
    LSG_LAUNCH_CHECK(ctx);
  }
  {
    ProfScope ps(ctx, prof_base >= 0 ? prof_base : PK_SCAN);
    k_radix_scan_tiles<<<R, 1024, 0, ctx->stream>>>(tile_hist, n_tiles, digit_total);
    LSG_LAUNCH_CHECK(ctx);
    k_radix_scan_digits<<<1, R, 0, ctx->stream>>>(digit_total, digit_base);
    LSG_LAUNCH_CHECK(ctx);
  }
  {
    ProfScope ps(ctx, prof_base >= 0 ? prof_base : PK_SCATTER);
    k_radix_scatter<KeyT, BITS><<<n_tiles, SORT_BLOCK, smem, ctx->stream>>>(kin, iin, kout, iout, n, shift, tile_hist, digit_base, n_tiles);
    LSG_LAUNCH_CHECK(ctx);
  }
  return 0;
}

// how the key's bits are shared between the passes: as few passes as 10-bit digits allow, digits of equal width (the
// last one takes what is left), never narrower than 8 bits unless the key itself is
inline int radix_plan(int bits, int *width) {
  if (bits < 1) bits = 1;
  const int passes = (bits + RADIX_MAX_BITS - 1) / RADIX_MAX_BITS;
  int w = (bits + passes - 1) / passes;
  if (w < 8) w = 8;
  *width = w;
  return (bits + w - 1) / w;
}

// LSD radix sort of (key, idx) pairs over key bits [0, bits).  Result pointers are returned (ping or pong).
template <typename KeyT>
int radix_sort_pairs(lsg_ctx *ctx, KeyT *keys_a, KeyT *keys_b, uint32_t *idx_a, uint32_t *idx_b, size_t n, int bits,
                     int slot_hist, KeyT **keys_out, uint32_t **idx_out, int prof_base = -1) {
  const uint32_t n_tiles = (uint32_t)((n + SORT_TILE - 1) / SORT_TILE);
  int width = 8;
  const int passes = radix_plan(bits, &width);
  const size_t R = (size_t)1 << width;
  void *hist_v = nullptr;
  LSG_TRY(lsg_ws(ctx, slot_hist, (R * n_tiles + 2 * R) * 4, &hist_v));
  uint32_t *tile_hist = (uint32_t *)hist_v;
  uint32_t *digit_total = tile_hist + R * n_tiles;
  uint32_t *digit_base = digit_total + R;
  KeyT *kin = keys_a, *kout = keys_b;
  uint32_t *iin = nullptr, *iout = idx_b;  // first pass: identity indices
  uint32_t *ia = idx_a, *ib = idx_b;
  for (int p = 0; p < passes; ++p) {
    const int shift = p * width;
    iout = (p & 1) ? ia : ib;
    if (width == 8) LSG_TRY((radix_pass<KeyT, 8>(ctx, kin, iin, kout, iout, n, shift, n_tiles, tile_hist, digit_total, digit_base, prof_base)));
    else if (width == 9) LSG_TRY((radix_pass<KeyT, 9>(ctx, kin, iin, kout, iout, n, shift, n_tiles, tile_hist, digit_total, digit_base, prof_base)));
    else LSG_TRY((radix_pass<KeyT, 10>(ctx, kin, iin, kout, iout, n, shift, n_tiles, tile_hist, digit_total, digit_base, prof_base)));
    KeyT *t = kin; kin = kout; kout = t;
    iin = iout;
  }
  *keys_out = kin;
  *idx_out = iin;
  return 0;
}

int bits_of(uint64_t v) {
  int b = 0;
  while (v) { b++; v >>= 1; }
  return b;
}

}  // namespace

// Sort events by (grp, quantised time).  Fills `out`; records land in `sorted_out` (or a workspace buffer if NULL).
int lsg_sort_events(lsg_ctx *ctx, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                    int64_t window_ns, bool exact, lsg_event *sorted_out, SortedEvents *out) {
  LSG_TRY(lsg_sort_events_prepare(ctx, events, n, idents, n_idents, window_ns, exact, out, nullptr, 0, nullptr));
  return lsg_sort_events_finish(ctx, events, n, idents, n_idents, sorted_out, out, !exact, true);
}

// phase 1: one pass over the events for the key layout (time range, group count) and the analytic counters
int lsg_sort_events_prepare(lsg_ctx *ctx, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                            int64_t window_ns, bool exact, SortedEvents *out, const lsg_span *spans, size_t n_spans,
                            const LsgKeyHint *hint) {
  memset(out, 0, sizeof *out);
  out->n = n;
  if (n >= ((size_t)1 << 32)) return lsg_fail(ctx, LSG_ERR_RANGE, "n_events %zu does not fit 32-bit indices", n);
  void *small_v = nullptr;
  LSG_TRY(lsg_ws(ctx, WS_SMALL, 4096, &small_v));
  EventStats *st_dev = (EventStats *)small_v;
  EventStats init;
  memset(&init, 0, sizeof init);
  init.ts_min = INT64_MAX;
  init.ts_max = INT64_MIN;
  EventStats *pin = (EventStats *)ctx->pinned;
  *pin = init;
  LSG_CUDA(ctx, cudaMemcpyAsync(st_dev, pin, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
  const bool from_spans = spans != nullptr && n_spans != 0;
  {
  ProfScope ps_stats(ctx, PK_STATS);
  if (n_idents) {
    k_ident_stats<<<grid_for(ctx, n_idents, 256, 1), 256, 0, ctx->stream>>>(idents, (uint32_t)n_idents, st_dev);
    LSG_LAUNCH_CHECK(ctx);
  }
  if (from_spans) {
    k_span_stats<<<grid_for(ctx, n_spans, 256, 8), 256, 0, ctx->stream>>>(spans, n_spans, idents, (uint32_t)n_idents, st_dev);
    LSG_LAUNCH_CHECK(ctx);
  } else if (n) {
    k_event_stats<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(events, n, idents, (uint32_t)n_idents, st_dev);
    LSG_LAUNCH_CHECK(ctx);
  }
  }
  EventStats st;
  if (from_spans && hint) {
    // no readback and no synchronisation: the key layout follows from the caller's statement
    k_check_hint<<<1, 1, 0, ctx->stream>>>(st_dev, hint->span_ts_lo, hint->span_ts_hi, hint->max_grp);
    LSG_LAUNCH_CHECK(ctx);
    st = init;
    st.ts_min = hint->span_ts_lo;
    st.ts_max = hint->span_ts_hi;
    st.max_grp = hint->max_grp;
  } else {
    LSG_CUDA(ctx, cudaMemcpyAsync(pin, st_dev, sizeof init, cudaMemcpyDeviceToHost, ctx->stream));
    LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    st = *pin;
  }
  if (st.bad & 1u) return lsg_fail(ctx, LSG_ERR_RANGE, "an event's ident is outside the identity table (n_idents=%zu)", n_idents);
  if (st.bad & 2u) return lsg_fail(ctx, LSG_ERR_RANGE, "a span's ident is outside the identity table (n_idents=%zu)", n_idents);
  out->stats_dev = st_dev;
  out->tally_in_sort = from_spans;
  KeySpec ks;
  uint64_t range = 0;
  if (from_spans) {
    // the events are not read before the sort: its first pass counts the unsupported types and checks the idents.
    // An event further than the window from every live span can match nothing, so the time axis is the spans'.
    out->n_live = n;
    if (st.ts_min <= st.ts_max) {
      ks.ts_min = st.ts_min < INT64_MIN + window_ns ? INT64_MIN : st.ts_min - window_ns;
      ks.ts_hi = st.ts_max > INT64_MAX - window_ns ? INT64_MAX : st.ts_max + window_ns;
      range = (uint64_t)ks.ts_hi - (uint64_t)ks.ts_min;
    } else {
      ks.ts_min = 0;    // no span can match: every event is dead
      ks.ts_hi = -1;
    }
  } else {
    out->n_live = st.n_live;
    ks.ts_min = st.n_live ? st.ts_min : 0;
    ks.ts_hi = st.n_live ? st.ts_max : -1;
    range = st.n_live ? (uint64_t)st.ts_max - (uint64_t)st.ts_min : 0;
  }
  ks.shift = 0;
  if (!exact) {
    // largest quantum <= window/32 (measured optimum on C2: halves the bin table for ~3% more inner-range events)
    while (ks.shift < 40 && ((int64_t)1 << (ks.shift + 1)) <= window_ns / 32) ks.shift++;
  }
  ks.tbits = (uint32_t)bits_of(range >> ks.shift);
  if (ks.tbits == 0) ks.tbits = 1;
  const int gbits = bits_of((uint64_t)st.max_grp + 1);   // one extra group id for the dead records
  ks.bits = ks.tbits + gbits;
  ks.dead = ((uint64_t)st.max_grp + 1) << ks.tbits;
  if (ks.bits > 62)
    return lsg_fail(ctx, LSG_ERR_RANGE, "sort key needs %u bits (time range %llu ns, %u groups): more than 62", ks.bits,
                    (unsigned long long)range, st.max_grp);
  out->key = ks;
  ctx->last_key = ks;
  return 0;
}

// phase 2: place the events.  Counting sort when the key space is small (and the caller does not need a stable,
// exact order), else keys + radix passes + gather.
int lsg_sort_events_finish(lsg_ctx *ctx, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                           lsg_event *sorted_out, SortedEvents *out, bool allow_counting, bool want_copy) {
  const KeySpec ks = out->key;
  if (n == 0) return 0;
  EventStats *tally_dst = out->tally_in_sort ? const_cast<EventStats *>(out->stats_dev) : nullptr;
  // where the events go: the caller's record array, or columns in the workspace (padded so that a 16-byte aligned
  // bulk copy may run a little past the last event)
  int64_t *c_ts = nullptr; double *c_val = nullptr; uint32_t *c_ident = nullptr, *c_pid = nullptr, *c_conn = nullptr, *c_trsig = nullptr;
  if (!sorted_out && want_copy) {
    const size_t n_pad = ((n + 3) & ~(size_t)3) + 4;
    void *sv;
    LSG_TRY(lsg_ws(ctx, WS_EV_SORTED, n_pad * 32, &sv));
    char *base = (char *)sv;
    c_ts = (int64_t *)base;
    c_val = (double *)(base + 8 * n_pad);
    c_ident = (uint32_t *)(base + 16 * n_pad); c_pid = (uint32_t *)(base + 20 * n_pad);
    c_conn = (uint32_t *)(base + 24 * n_pad); c_trsig = (uint32_t *)(base + 28 * n_pad);
    out->col = EventColumns{c_ts, c_val, c_ident, c_pid, c_conn, c_trsig};
  } else {
    out->rec = sorted_out;
  }
  auto gather = [&](const uint32_t *perm, size_t count, const uint32_t *n_placed) -> int {
    if (!sorted_out && !want_copy) return 0;   // the consumer reads the input records through the permutation
    ProfScope ps(ctx, PK_GATHER);
    if (sorted_out) k_gather_events<<<grid_for(ctx, count, 256, 8), 256, 0, ctx->stream>>>(events, perm, count, n_placed, sorted_out);
    else k_gather_events_soa<<<grid_for(ctx, count, 256, 8), 256, 0, ctx->stream>>>(events, perm, count, n_placed, c_ts, c_val, c_ident, c_pid, c_conn, c_trsig);
    LSG_LAUNCH_CHECK(ctx);
    return 0;
  };
  static const uint64_t counting_max = getenv("LSG_COUNTING_MAX_BITS") ? (uint64_t)1 << atoi(getenv("LSG_COUNTING_MAX_BITS")) : COUNTING_MAX_KEYS;
  if (allow_counting && ks.shift != 0 && ks.dead <= counting_max) {
    const size_t n_keys = (size_t)ks.dead + 1;   // live keys are < dead; one extra entry so that bins[k + 1] always exists
    void *bv, *sv, *pv, *kv;
    LSG_TRY(lsg_ws(ctx, WS_BINS, (n_keys + SCAN_TILE) * 4, &bv));
    const uint32_t n_blocks = (uint32_t)((n_keys + SCAN_TILE - 1) / SCAN_TILE);
    LSG_TRY(lsg_ws(ctx, WS_SCAN, (size_t)n_blocks * 4 + 64, &sv));
    LSG_TRY(lsg_ws(ctx, WS_IDX_A, n * 4 + 64, &pv));
    LSG_TRY(lsg_ws(ctx, WS_KEYS_A, n * 8, &kv));
    uint32_t *bins = (uint32_t *)bv, *sums = (uint32_t *)sv, *perm = (uint32_t *)pv;
    uint2 *key_rank = (uint2 *)kv;
    {
      ProfScope ps(ctx, PK_HIST);
      LSG_CUDA(ctx, cudaMemsetAsync(bins, 0, n_keys * 4, ctx->stream));
      k_count_keys<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(events, n, idents, (uint32_t)n_idents, ks, bins, key_rank,
                                                                      tally_dst);
      LSG_LAUNCH_CHECK(ctx);
    }
    {
      ProfScope ps(ctx, PK_SCAN);
      k_scan_reduce<<<n_blocks, SCAN_BLOCK, 0, ctx->stream>>>(bins, n_keys, sums);
      LSG_LAUNCH_CHECK(ctx);
      k_scan_sums<<<1, 1024, 0, ctx->stream>>>(sums, n_blocks);
      LSG_LAUNCH_CHECK(ctx);
      k_scan_apply<<<n_blocks, SCAN_BLOCK, 0, ctx->stream>>>(bins, n_keys, sums);   // bins[k] = start of key k's run
      LSG_LAUNCH_CHECK(ctx);
    }
    {
      ProfScope ps(ctx, PK_SCATTER);
      k_place_events<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(key_rank, n, bins, perm);
      LSG_LAUNCH_CHECK(ctx);
    }
    // after the scan bins[dead] is the start of the (empty) dead run: the number of events that were placed
    if (out->n_live) LSG_TRY(gather(perm, out->n_live, bins + ks.dead));
    out->bins = bins;
    out->perm = perm;
    return 0;
  }
  const bool k32 = ks.bits <= 31;
  void *ka, *kb, *ia, *ib;
  LSG_TRY(lsg_ws(ctx, WS_KEYS_A, n * (k32 ? 4 : 8), &ka));
  LSG_TRY(lsg_ws(ctx, WS_KEYS_B, n * (k32 ? 4 : 8), &kb));
  LSG_TRY(lsg_ws(ctx, WS_IDX_A, n * 4 + 64, &ia));
  LSG_TRY(lsg_ws(ctx, WS_IDX_B, n * 4 + 64, &ib));
  uint32_t *perm = nullptr;
  if (k32) {
    {
      ProfScope ps(ctx, PK_KEYS);
      k_event_keys<uint32_t><<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(events, n, idents, (uint32_t)n_idents, ks, (uint32_t *)ka, tally_dst);
      LSG_LAUNCH_CHECK(ctx);
    }
    uint32_t *ko;
    LSG_TRY(radix_sort_pairs<uint32_t>(ctx, (uint32_t *)ka, (uint32_t *)kb, (uint32_t *)ia, (uint32_t *)ib, n, ks.bits, WS_HIST, &ko, &perm));
    out->keys = ko;
  } else {
    {
      ProfScope ps(ctx, PK_KEYS);
      k_event_keys<uint64_t><<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(events, n, idents, (uint32_t)n_idents, ks, (uint64_t *)ka, tally_dst);
      LSG_LAUNCH_CHECK(ctx);
    }
    uint64_t *ko;
    LSG_TRY(radix_sort_pairs<uint64_t>(ctx, (uint64_t *)ka, (uint64_t *)kb, (uint32_t *)ia, (uint32_t *)ib, n, ks.bits, WS_HIST, &ko, &perm));
    out->keys = ko;
  }
  out->perm = perm;
  return gather(perm, n, nullptr);
}

// Processing order of the spans: same key as the events, so consecutive spans read overlapping event ranges.
int lsg_sort_spans(lsg_ctx *ctx, const lsg_span *spans, size_t n, const lsg_ident *idents, size_t n_idents, const KeySpec &ks,
                   const uint32_t **perm_out) {
  *perm_out = nullptr;
  if (n == 0) return 0;
  if (n >= ((size_t)1 << 32)) return lsg_fail(ctx, LSG_ERR_RANGE, "n_spans %zu does not fit 32-bit indices", n);
  const bool k32 = ks.bits <= 31;
  void *ka, *kb, *ia, *ib;
  LSG_TRY(lsg_ws(ctx, WS_SPAN_KEYS_A, n * (k32 ? 4 : 8), &ka));
  LSG_TRY(lsg_ws(ctx, WS_SPAN_KEYS_B, n * (k32 ? 4 : 8), &kb));
  LSG_TRY(lsg_ws(ctx, WS_SPAN_IDX_A, n * 4, &ia));
  LSG_TRY(lsg_ws(ctx, WS_SPAN_IDX_B, n * 4, &ib));
  // ts_max of the key space: the largest quantised time representable
  const int64_t ts_max = (int64_t)((uint64_t)ks.ts_min + ((((uint64_t)1 << ks.tbits) - 1) << ks.shift));
  uint32_t *perm = nullptr;
  ProfScope ps_span(ctx, PK_SPAN_SORT);
  const bool saved_prof = ctx->prof;
  ctx->prof = false;   // the passes of the span sort are accounted as one PK_SPAN_SORT interval
  struct Restore { lsg_ctx *c; bool v; ~Restore() { c->prof = v; } } restore{ctx, saved_prof};
  if (k32) {
    k_span_keys<uint32_t><<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(spans, n, idents, (uint32_t)n_idents, ks, ts_max, (uint32_t *)ka);
    LSG_LAUNCH_CHECK(ctx);
    uint32_t *ko;
    LSG_TRY(radix_sort_pairs<uint32_t>(ctx, (uint32_t *)ka, (uint32_t *)kb, (uint32_t *)ia, (uint32_t *)ib, n, ks.bits, WS_SPAN_HIST, &ko, &perm));
  } else {
    k_span_keys<uint64_t><<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(spans, n, idents, (uint32_t)n_idents, ks, ts_max, (uint64_t *)ka);
    LSG_LAUNCH_CHECK(ctx);
    uint64_t *ko;
    LSG_TRY(radix_sort_pairs<uint64_t>(ctx, (uint64_t *)ka, (uint64_t *)kb, (uint32_t *)ia, (uint32_t *)ib, n, ks.bits, WS_SPAN_HIST, &ko, &perm));
  }
  *perm_out = perm;
  return 0;
}

// ---- bucket partition: the records themselves, once, into coarse (group, time slab) buckets ----------------------
// The permutation paths above leave the 32-byte records where they are and let the join gather them (random 32-byte
// reads: ~113 B of DRAM traffic each, and the join waits on their latency).  Here the records are MOVED, once, into
// buckets of one join group x 64 time quanta (a little longer than the window):
//   k_bucket_count    one sequential read of the records; one no-return atomic (RED) per record on its bucket's counter
//   k_scan_*          counts -> bucket starts
//   k_bucket_scatter  second sequential read; slot = returning atomic on the bucket's cursor; one 256-bit store of the
//                     record and a 4-byte store of its input index.  Records of one bucket that are in flight together
//                     land in neighbouring slots, so the partially written 128-byte lines sit in L2 until they are
//                     complete: DRAM sees whole lines.  That holds while a line's four arrivals are at most a few
//                     million records apart, i.e. up to ~10^5 buckets; beyond that (1 B events: millions of buckets)
//                     the scatter runs twice, first into <= 4096 super-buckets (runs of consecutive buckets), then,
//                     reading those sequentially, into the buckets themselves.
// Records that can match nothing (unsupported type, zero time, group 0, outside the spans' time range) are dropped
// (events) or collected behind the last bucket (spans: they still get their empty outputs).
namespace {

struct EventBucketFn {
  const lsg_ident *idents;
  uint32_t n_idents;
  KeySpec ks;
  __device__ __forceinline__ bool operator()(const U8 &u, EventTally &t, uint32_t *b) const {
    EventRec e;
    e.ts = (int64_t)(((uint64_t)u.a.y << 32) | u.a.x);
    e.ident = u.b.x; e.pid = u.b.y; e.conn = u.b.z; e.trace_sig = u.b.w;
    uint64_t k;
    if (!event_key(e, idents, n_idents, ks, t, &k)) return false;
    *b = (uint32_t)(k >> LSG_BUCKET_SHIFT);
    return true;
  }
};

struct SpanBucketFn {
  const lsg_ident *idents;
  uint32_t n_idents;
  KeySpec ks;
  uint32_t n_buckets;
  __device__ __forceinline__ bool operator()(const U8 &u, EventTally &, uint32_t *b) const {
    const int64_t ts = (int64_t)(((uint64_t)u.a.y << 32) | u.a.x);
    const uint32_t ident = u.a.z;
    *b = n_buckets;   // cannot match: behind every real bucket (a bad ident is reported by k_span_stats)
    const uint32_t g = ident < n_idents ? load_ident(idents, ident).grp : 0u;
    if (g != 0 && ts != LSG_TS_ZERO && ts >= ks.ts_min && ts <= ks.ts_hi) {
      const uint64_t k = lsg_make_key(ks, g, ts);
      if (k < ks.dead) *b = (uint32_t)(k >> LSG_BUCKET_SHIFT);
    }
    return true;
  }
};

template <typename Fn>
__global__ void __launch_bounds__(256)
k_bucket_count(const lsg_event *__restrict__ recs, size_t n, const Fn fn, uint32_t *__restrict__ counts, EventStats *st) {
  EventTally tally;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    U8 r[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) r[u] = ld_nc_v8(recs + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      uint32_t b;
      if (fn(r[u], tally, &b)) atomicAdd(&counts[b], 1u);   // result unused: RED
    }
  }
  for (; i < n; i += stride) {
    uint32_t b;
    if (fn(ld_nc_v8(recs + i), tally, &b)) atomicAdd(&counts[b], 1u);
  }
  tally.flush(st);
}

// cursor[b >> sshift] is advanced; with sshift == 0 the cursors are the bucket starts themselves and end up as the
// bucket ends.  orig_in NULL: the records are the caller's input (index = position).
template <typename Fn>
__global__ void __launch_bounds__(256)
k_bucket_scatter(const lsg_event *__restrict__ recs, const uint32_t *__restrict__ orig_in, size_t n, const Fn fn,
                 uint32_t sshift, uint32_t *__restrict__ cursor, lsg_event *__restrict__ out, uint32_t *__restrict__ orig_out) {
  EventTally tally;   // not flushed: the count pass already told
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    U8 r[4];
    uint32_t b[4], slot[4], src[4];
    bool ok[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) r[u] = ld_nc_v8(recs + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) src[u] = orig_in ? __ldg(orig_in + i + u * stride) : (uint32_t)(i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) ok[u] = fn(r[u], tally, &b[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u) slot[u] = ok[u] ? atomicAdd(&cursor[b[u] >> sshift], 1u) : 0u;   // four round trips in flight
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (ok[u]) {
        st_v8(out + slot[u], r[u]);
        orig_out[slot[u]] = src[u];
      }
  }
  for (; i < n; i += stride) {
    const U8 r = ld_nc_v8(recs + i);
    uint32_t b;
    if (fn(r, tally, &b)) {
      const uint32_t slot = atomicAdd(&cursor[b >> sshift], 1u);
      st_v8(out + slot, r);
      orig_out[slot] = orig_in ? orig_in[i] : (uint32_t)i;
    }
  }
}

// second level of the two-level partition: the input is the first level's output, whose length is only known on the
// device (*n_dev: the last super-bucket's cursor)
template <typename Fn>
__global__ void __launch_bounds__(256)
k_bucket_scatter_tail(const lsg_event *__restrict__ recs, const uint32_t *__restrict__ orig_in, const uint32_t *__restrict__ n_dev,
                      const Fn fn, uint32_t *__restrict__ cursor, lsg_event *__restrict__ out, uint32_t *__restrict__ orig_out) {
  EventTally tally;
  const size_t n = *n_dev;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    U8 r[4];
    uint32_t b[4], slot[4], src[4];
    bool ok[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) r[u] = ld_nc_v8(recs + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) src[u] = __ldg(orig_in + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u) ok[u] = fn(r[u], tally, &b[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u) slot[u] = ok[u] ? atomicAdd(&cursor[b[u]], 1u) : 0u;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (ok[u]) {
        st_v8(out + slot[u], r[u]);
        orig_out[slot[u]] = src[u];
      }
  }
  for (; i < n; i += stride) {
    const U8 r = ld_nc_v8(recs + i);
    uint32_t b;
    if (fn(r, tally, &b)) {
      const uint32_t slot = atomicAdd(&cursor[b], 1u);
      st_v8(out + slot, r);
      orig_out[slot] = orig_in[i];
    }
  }
}

__global__ void k_super_starts(const uint32_t *__restrict__ starts, uint32_t n_super, uint32_t sshift, uint32_t *__restrict__ cur) {
  for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n_super; s += gridDim.x * blockDim.x)
    cur[s] = starts[(size_t)s << sshift];
}

constexpr uint32_t BK_SINGLE_LEVEL_MAX = 1u << 17;   // buckets a single scatter keeps line-local in L2
constexpr uint32_t BK_SUPER_BITS = 12;               // at most 4096 super-buckets

template <typename Fn>
int bucket_partition(lsg_ctx *ctx, const void *recs, size_t n, const Fn &fn, uint32_t n_slots /* counters */,
                     EventStats *stats_dev, int ws_rec, int ws_orig, int ws_ends, int ws_sums, int ws_tmp, int pk_count,
                     int pk_scatter, Bucketed *out) {
  // layout of the counter array: [0..3] zero (ends[-1] = 0), [4 .. 4 + n_slots) counts -> starts -> ends
  void *ev, *ov, *cv, *sv;
  LSG_TRY(lsg_ws(ctx, ws_rec, (n + 1) * 32, &ev));
  LSG_TRY(lsg_ws(ctx, ws_orig, (n + 1) * 4, &ov));
  LSG_TRY(lsg_ws(ctx, ws_ends, ((size_t)n_slots + 8 + SCAN_TILE) * 4, &cv));
  const uint32_t n_blocks = (uint32_t)(((size_t)n_slots + SCAN_TILE - 1) / SCAN_TILE);
  LSG_TRY(lsg_ws(ctx, ws_sums, (size_t)n_blocks * 4 + 64, &sv));   // (events and spans are partitioned on two streams at once)
  uint32_t *arr = (uint32_t *)cv, *cnt = arr + 4, *sums = (uint32_t *)sv;
  const lsg_event *in = (const lsg_event *)recs;
  {
    ProfScope ps(ctx, pk_count);
    LSG_CUDA(ctx, cudaMemsetAsync(arr, 0, ((size_t)n_slots + 8) * 4, ctx->stream));
    if (n) {
      k_bucket_count<Fn><<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(in, n, fn, cnt, stats_dev);
      LSG_LAUNCH_CHECK(ctx);
    }
    k_scan_reduce<<<n_blocks, SCAN_BLOCK, 0, ctx->stream>>>(cnt, n_slots, sums);
    LSG_LAUNCH_CHECK(ctx);
    k_scan_sums<<<1, 1024, 0, ctx->stream>>>(sums, n_blocks);
    LSG_LAUNCH_CHECK(ctx);
    k_scan_apply<<<n_blocks, SCAN_BLOCK, 0, ctx->stream>>>(cnt, n_slots, sums);   // cnt[b] = start of bucket b
    LSG_LAUNCH_CHECK(ctx);
  }
  if (n) {
    ProfScope ps(ctx, pk_scatter);
    if (n_slots <= BK_SINGLE_LEVEL_MAX) {
      k_bucket_scatter<Fn><<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(in, nullptr, n, fn, 0u, cnt, (lsg_event *)ev,
                                                                             (uint32_t *)ov);
      LSG_LAUNCH_CHECK(ctx);
    } else {
      uint32_t sshift = 0;
      while (((n_slots - 1) >> sshift) >= (1u << BK_SUPER_BITS)) sshift++;
      const uint32_t n_super = ((n_slots - 1) >> sshift) + 1;
      void *tv, *tov, *suv;
      LSG_TRY(lsg_ws(ctx, ws_tmp, (n + 1) * 32, &tv));       // three consecutive slots: records, indices, cursors
      LSG_TRY(lsg_ws(ctx, ws_tmp + 1, (n + 1) * 4, &tov));
      LSG_TRY(lsg_ws(ctx, ws_tmp + 2, (size_t)n_super * 4 + 64, &suv));
      k_super_starts<<<grid_for(ctx, n_super, 256, 1), 256, 0, ctx->stream>>>(cnt, n_super, sshift, (uint32_t *)suv);
      LSG_LAUNCH_CHECK(ctx);
      k_bucket_scatter<Fn><<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(in, nullptr, n, fn, sshift, (uint32_t *)suv,
                                                                             (lsg_event *)tv, (uint32_t *)tov);
      LSG_LAUNCH_CHECK(ctx);
      // everything that was placed is now grouped by super-bucket in [0, n_placed); n_placed <= n is not known on the
      // host, so the second pass reads a device-side count: the cursor of the last super-bucket is the total
      // (the tail [n_placed, n) of tmp is stale: the second pass must not read it)
      k_bucket_scatter_tail<Fn><<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>((const lsg_event *)tv, (const uint32_t *)tov,
                                                                                  (const uint32_t *)suv + (n_super - 1), fn, cnt,
                                                                                  (lsg_event *)ev, (uint32_t *)ov);
      LSG_LAUNCH_CHECK(ctx);
    }
  }
  out->rec = ev;
  out->orig = (const uint32_t *)ov;
  out->ends = cnt;
  return 0;
}

}  // namespace

KeySpec lsg_bucket_keyspec(const KeySpec &ks_in, size_t n_events, int64_t window_ns) {
  KeySpec ks = ks_in;
  const uint64_t max_grp = (ks.dead >> ks.tbits) - 1;
  const uint64_t range = ks.ts_hi >= ks.ts_min ? (uint64_t)ks.ts_hi - (uint64_t)ks.ts_min : 0;
  auto layout = [&]() {
    ks.tbits = (uint32_t)bits_of(range >> ks.shift);
    if (ks.tbits < LSG_BUCKET_SHIFT) ks.tbits = LSG_BUCKET_SHIFT;   // a bucket never straddles two groups
    ks.bits = ks.tbits + (uint32_t)bits_of(max_grp + 1);
    ks.dead = (max_grp + 1) << ks.tbits;
  };
  layout();
  // a sparse batch (long time range, few events) would get millions of nearly empty buckets: coarsen the quantum
  // until a bucket holds ~16 events on average (ranges stay supersets of the exact windows, the pairs are re-checked)
  (void)window_ns;
  while (ks.shift < 40 && (ks.dead >> LSG_BUCKET_SHIFT) * 16 > (uint64_t)n_events + 1024 && ks.tbits > LSG_BUCKET_SHIFT) {
    ks.shift++;
    layout();
  }
  return ks;
}

int lsg_bucket_events(lsg_ctx *ctx, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                      const KeySpec &bk, LsgBatchStats *stats_dev, Bucketed *out) {
  memset(out, 0, sizeof *out);
  out->key = bk;
  out->n_buckets = (uint32_t)(bk.dead >> LSG_BUCKET_SHIFT);
  EventBucketFn fn{idents, (uint32_t)n_idents, bk};
  return bucket_partition(ctx, events, n, fn, out->n_buckets, stats_dev, WS_BK_EV, WS_BK_EV_ORIG, WS_BK_EV_ENDS, WS_SCAN, WS_BK_TMP, PK_BK_COUNT,
                          PK_BK_SCATTER, out);
}

int lsg_bucket_spans(lsg_ctx *ctx, const lsg_span *spans, size_t n, const lsg_ident *idents, size_t n_idents, const KeySpec &bk,
                     Bucketed *out) {
  memset(out, 0, sizeof *out);
  out->key = bk;
  out->n_buckets = (uint32_t)(bk.dead >> LSG_BUCKET_SHIFT);
  SpanBucketFn fn{idents, (uint32_t)n_idents, bk, out->n_buckets};
  return bucket_partition(ctx, spans, n, fn, out->n_buckets + 1, nullptr, WS_BK_SP, WS_BK_SP_ORIG, WS_BK_SP_ENDS, WS_SPAN_HIST, WS_SPAN_KEYS_A, PK_BK_SPANS,
                          PK_BK_SPANS, out);
}

// ---- F1 support: the events of one signal in stable (pod, exact time) order ---------------------------------
namespace {
struct PodStats { long long ts_min, ts_max; unsigned long long n_live; unsigned int max_pod, bad; unsigned int pad[2]; };

__device__ __forceinline__ bool podtime_live(const EventRec &e, const lsg_ident *idents, uint32_t n_idents, uint32_t sig,
                                             uint32_t *pod, bool *bad) {
  *pod = 0;
  if (e.ident >= n_idents) { *bad = true; return false; }
  if ((e.trace_sig & 15u) != sig || e.ts == LSG_TS_ZERO) return false;
  *pod = load_ident(idents, e.ident).pod;
  return true;   // the reference keys its buckets by the pod string, empty or not (retry_storm.go:50-55)
}

__global__ void __launch_bounds__(256)
k_podtime_stats(const lsg_event *__restrict__ events, size_t n, const lsg_ident *__restrict__ idents, uint32_t n_idents,
                uint32_t sig, PodStats *st) {
  long long tmin = INT64_MAX, tmax = INT64_MIN;
  unsigned int live = 0, mp = 0;
  bool bad = false;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const EventRec e = load_event(events + i);
    uint32_t pod;
    if (podtime_live(e, idents, n_idents, sig, &pod, &bad)) {
      live++;
      mp = max(mp, pod);
      tmin = min(tmin, (long long)e.ts);
      tmax = max(tmax, (long long)e.ts);
    }
  }
  for (int o = 16; o; o >>= 1) {
    tmin = min(tmin, __shfl_xor_sync(0xffffffffu, tmin, o));
    tmax = max(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
    live += __shfl_xor_sync(0xffffffffu, live, o);
    mp = max(mp, __shfl_xor_sync(0xffffffffu, mp, o));
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(&st->bad, 1u);
  if ((threadIdx.x & 31) == 0 && live) {
    atomicMin(&st->ts_min, tmin);
    atomicMax(&st->ts_max, tmax);
    atomicAdd(&st->n_live, (unsigned long long)live);
    atomicMax(&st->max_pod, mp);
  }
}

__global__ void __launch_bounds__(256)
k_podtime_keys(const lsg_event *__restrict__ events, size_t n, const lsg_ident *__restrict__ idents, uint32_t n_idents,
               uint32_t sig, KeySpec ks, uint64_t *__restrict__ keys) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const EventRec e = load_event(events + i);
    uint32_t pod;
    bool bad = false;
    uint64_t k = ks.dead;
    if (podtime_live(e, idents, n_idents, sig, &pod, &bad)) k = lsg_make_key(ks, pod, e.ts);
    keys[i] = k;
  }
}
}  // namespace

int lsg_sort_signal_by_pod(lsg_ctx *ctx, const lsg_event *events, size_t n, const lsg_ident *idents, size_t n_idents,
                           uint32_t sig, const uint64_t **keys_out, const uint32_t **perm_out, KeySpec *ks_out,
                           size_t *n_live_out) {
  *keys_out = nullptr; *perm_out = nullptr; *n_live_out = 0;
  if (n >= ((size_t)1 << 32)) return lsg_fail(ctx, LSG_ERR_RANGE, "n_events %zu does not fit 32-bit indices", n);
  void *small_v = nullptr;
  LSG_TRY(lsg_ws(ctx, WS_SMALL, 4096, &small_v));
  PodStats *st_dev = (PodStats *)small_v;
  PodStats *pin = (PodStats *)ctx->pinned;
  memset(pin, 0, sizeof *pin);
  pin->ts_min = INT64_MAX;
  pin->ts_max = INT64_MIN;
  LSG_CUDA(ctx, cudaMemcpyAsync(st_dev, pin, sizeof *pin, cudaMemcpyHostToDevice, ctx->stream));
  if (n) {
    k_podtime_stats<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(events, n, idents, (uint32_t)n_idents, sig, st_dev);
    LSG_LAUNCH_CHECK(ctx);
  }
  LSG_CUDA(ctx, cudaMemcpyAsync(pin, st_dev, sizeof *pin, cudaMemcpyDeviceToHost, ctx->stream));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const PodStats st = *pin;
  if (st.bad) return lsg_fail(ctx, LSG_ERR_RANGE, "an event's ident is outside the identity table (n_idents=%zu)", n_idents);
  KeySpec ks;
  ks.ts_min = st.n_live ? st.ts_min : 0;
  ks.shift = 0;
  ks.tbits = (uint32_t)bits_of(st.n_live ? (uint64_t)st.ts_max - (uint64_t)st.ts_min : 0);
  if (ks.tbits == 0) ks.tbits = 1;
  ks.bits = ks.tbits + bits_of((uint64_t)st.max_pod + 1);
  ks.dead = ((uint64_t)st.max_pod + 1) << ks.tbits;
  if (ks.bits > 62) return lsg_fail(ctx, LSG_ERR_RANGE, "(pod, time) key needs %u bits: more than 62", ks.bits);
  *ks_out = ks;
  *n_live_out = st.n_live;
  if (n == 0) return 0;
  void *ka, *kb, *ia, *ib;
  LSG_TRY(lsg_ws(ctx, WS_KEYS_A, n * 8, &ka));
  LSG_TRY(lsg_ws(ctx, WS_KEYS_B, n * 8, &kb));
  LSG_TRY(lsg_ws(ctx, WS_IDX_A, n * 4 + 64, &ia));
  LSG_TRY(lsg_ws(ctx, WS_IDX_B, n * 4 + 64, &ib));
  k_podtime_keys<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(events, n, idents, (uint32_t)n_idents, sig, ks, (uint64_t *)ka);
  LSG_LAUNCH_CHECK(ctx);
  uint64_t *ko;
  uint32_t *perm;
  LSG_TRY(radix_sort_pairs<uint64_t>(ctx, (uint64_t *)ka, (uint64_t *)kb, (uint32_t *)ia, (uint32_t *)ib, n, ks.bits, WS_HIST, &ko, &perm));
  *keys_out = ko;
  *perm_out = perm;
  return 0;
}

// ---- F4: exact quantiles per segment -----------------------------------------------------------------------
// quantile() of pkg/slo/calculator.go:121-149 (the same body as pkg/releasegate/gate.go:737-760) over many value
// lists at once: a stable LSD sort by value, then by segment, leaves every segment's values contiguous and
// ascending; one thread per (segment, q) then reads the two neighbours of pos = q * (count - 1).
namespace {
constexpr int QH_MAX = 4096;   // segment histograms up to this size are accumulated in shared memory first

// sort.Float64s puts NaN before every number; f64_key() of a non-NaN is never 0
__device__ __forceinline__ unsigned long long quant_key(double v) { return isnan(v) ? 0ull : f64_key(v); }

__global__ void __launch_bounds__(256)
k_quant_keys(const double *__restrict__ values, const uint32_t *__restrict__ seg, size_t n, uint32_t n_seg,
             uint64_t *__restrict__ keys, uint32_t *__restrict__ counts) {
  __shared__ uint32_t s_hist[QH_MAX];
  const bool local = n_seg + 1 <= (uint32_t)QH_MAX;
  if (local) {
    for (uint32_t k = threadIdx.x; k <= n_seg; k += blockDim.x) s_hist[k] = 0;
    __syncthreads();
  }
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    keys[i] = quant_key(values[i]);
    uint32_t s = seg ? seg[i] : 0u;
    if (s > n_seg) s = n_seg;             // out-of-range segment ids share one trailing bucket that nobody reads
    atomicAdd(local ? &s_hist[s] : &counts[s], 1u);
  }
  if (local) {
    __syncthreads();
    for (uint32_t k = threadIdx.x; k <= n_seg; k += blockDim.x)
      if (s_hist[k]) atomicAdd(&counts[k], s_hist[k]);
  }
}

__global__ void __launch_bounds__(256)
k_quant_segkeys(const uint32_t *__restrict__ seg, const uint32_t *__restrict__ perm, size_t n, uint32_t n_seg,
                uint32_t *__restrict__ segkeys) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    segkeys[i] = min(seg[perm[i]], n_seg);
}

// offsets[s] = start of segment s in the final order (exclusive scan of the counts); by_value = keys after the
// value sort; perm2 (NULL = identity) maps a final position to a position of by_value
__global__ void __launch_bounds__(256)
k_quant_pick(const uint64_t *__restrict__ by_value, const uint32_t *__restrict__ perm2, const uint32_t *__restrict__ offsets,
             uint32_t n_seg, const double *__restrict__ q, uint32_t n_q, double *__restrict__ out,
             uint64_t *__restrict__ count_out) {
  const size_t total = (size_t)n_seg * n_q;
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
    const uint32_t s = (uint32_t)(t / n_q), j = (uint32_t)(t % n_q);
    const uint32_t off = offsets[s], cnt = offsets[s + 1] - off;
    if (j == 0 && count_out) count_out[s] = cnt;
    if (!out) continue;                                      // counts only
    double r = 0.0;                                          // calculator.go:122-124
    if (cnt) {
      double qq = q[j];
      if (qq <= 0) qq = 0;                                   // calculator.go:125-130
      if (qq >= 1) qq = 1;
      auto at = [&](uint32_t k) { const uint32_t p = off + k; return key_f64(by_value[perm2 ? perm2[p] : p]); };
      if (cnt == 1) {
        r = at(0);                                           // calculator.go:136-138
      } else {
        const double pos = qq * (double)(cnt - 1);           // calculator.go:140-148
        const double lo = floor(pos), hi = ceil(pos);
        if (lo == hi) {
          r = at((uint32_t)lo);
        } else {
          const double frac = pos - lo;
          r = at((uint32_t)lo) * (1 - frac) + at((uint32_t)hi) * frac;
        }
      }
    }
    out[t] = r;
  }
}

template <typename T>
int quant_stage(lsg_ctx *ctx, int slot, const T *host, size_t n, const T **dev) {
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

int lsg_segment_quantiles_dev(lsg_ctx *ctx, const double *values, const uint32_t *segment, size_t n, size_t n_segments,
                              const double *q, size_t n_q, double *out, uint64_t *count_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if ((n && !values) || (n_q && !q) || (n_segments && n_q && !out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  if (n >= ((size_t)1 << 32)) return lsg_fail(ctx, LSG_ERR_RANGE, "n %zu does not fit 32-bit indices", n);
  if (n_segments >= ((size_t)1 << 31)) return lsg_fail(ctx, LSG_ERR_RANGE, "n_segments %zu is out of range", n_segments);
  if (n && !segment && n_segments > 1) return lsg_fail(ctx, LSG_ERR_INVALID, "segment is NULL but n_segments is %zu", n_segments);
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  if (n_segments == 0 || (n_q == 0 && !count_out)) return 0;
  const uint32_t n_seg = (uint32_t)n_segments;
  // counts[0..n_seg] (+ one trailing zero so that the scan yields n_seg + 2 offsets)
  const size_t n_cnt = (size_t)n_seg + 2;
  const uint32_t n_blocks = (uint32_t)((n_cnt + SCAN_TILE - 1) / SCAN_TILE);
  void *cnt_v, *sums_v;
  LSG_TRY(lsg_ws(ctx, WS_BINS, n_cnt * 4 + 64, &cnt_v));
  LSG_TRY(lsg_ws(ctx, WS_SCAN, (size_t)n_blocks * 4 + 64, &sums_v));
  uint32_t *counts = (uint32_t *)cnt_v;
  LSG_CUDA(ctx, cudaMemsetAsync(counts, 0, n_cnt * 4, ctx->stream));
  uint64_t *by_value = nullptr;
  uint32_t *perm2 = nullptr;
  if (n) {
    void *ka, *kb, *ia, *ib;
    LSG_TRY(lsg_ws(ctx, WS_KEYS_A, n * 8, &ka));
    LSG_TRY(lsg_ws(ctx, WS_KEYS_B, n * 8, &kb));
    LSG_TRY(lsg_ws(ctx, WS_IDX_A, n * 4 + 64, &ia));
    LSG_TRY(lsg_ws(ctx, WS_IDX_B, n * 4 + 64, &ib));
    k_quant_keys<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(values, segment, n, n_seg, (uint64_t *)ka, counts);
    LSG_LAUNCH_CHECK(ctx);
    uint32_t *perm1;
    LSG_TRY(radix_sort_pairs<uint64_t>(ctx, (uint64_t *)ka, (uint64_t *)kb, (uint32_t *)ia, (uint32_t *)ib, n, 64, WS_HIST,
                                       &by_value, &perm1));
    if (segment) {   // a stable second sort by segment keeps each segment's values ascending
      void *sa, *sb, *ja, *jb;
      LSG_TRY(lsg_ws(ctx, WS_SPAN_KEYS_A, n * 4 + 64, &sa));
      LSG_TRY(lsg_ws(ctx, WS_SPAN_KEYS_B, n * 4 + 64, &sb));
      LSG_TRY(lsg_ws(ctx, WS_SPAN_IDX_A, n * 4 + 64, &ja));
      LSG_TRY(lsg_ws(ctx, WS_SPAN_IDX_B, n * 4 + 64, &jb));
      k_quant_segkeys<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(segment, perm1, n, n_seg, (uint32_t *)sa);
      LSG_LAUNCH_CHECK(ctx);
      uint32_t *seg_sorted;
      LSG_TRY(radix_sort_pairs<uint32_t>(ctx, (uint32_t *)sa, (uint32_t *)sb, (uint32_t *)ja, (uint32_t *)jb, n,
                                         bits_of(n_seg), WS_SPAN_HIST, &seg_sorted, &perm2));
    }
  }
  k_scan_reduce<<<n_blocks, SCAN_BLOCK, 0, ctx->stream>>>(counts, n_cnt, (uint32_t *)sums_v);
  LSG_LAUNCH_CHECK(ctx);
  k_scan_sums<<<1, 1024, 0, ctx->stream>>>((uint32_t *)sums_v, n_blocks);
  LSG_LAUNCH_CHECK(ctx);
  k_scan_apply<<<n_blocks, SCAN_BLOCK, 0, ctx->stream>>>(counts, n_cnt, (uint32_t *)sums_v);
  LSG_LAUNCH_CHECK(ctx);
  const uint32_t nq_eff = n_q ? (uint32_t)n_q : 1u;   // n_q == 0: one thread per segment writes count_out only
  k_quant_pick<<<grid_for(ctx, (size_t)n_seg * nq_eff, 256, 8), 256, 0, ctx->stream>>>(by_value, perm2, counts, n_seg, q, nq_eff,
                                                                                      n_q ? out : nullptr, count_out);
  LSG_LAUNCH_CHECK(ctx);
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

int lsg_segment_quantiles(lsg_ctx *ctx, const double *values, const uint32_t *segment, size_t n, size_t n_segments,
                          const double *q, size_t n_q, double *out, uint64_t *count_out) {
  if (!ctx) return LSG_ERR_INVALID;
  if ((n && !values) || (n_q && !q) || (n_segments && n_q && !out)) return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  for (size_t j = 0; j < n_q; ++j)
    if (q[j] != q[j]) return lsg_fail(ctx, LSG_ERR_INVALID, "q[%zu] is NaN", j);
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  const double *d_val, *d_q;
  const uint32_t *d_seg;
  LSG_TRY(quant_stage(ctx, WS_IN0, values, n, &d_val));
  LSG_TRY(quant_stage(ctx, WS_IN1, segment, n, &d_seg));
  LSG_TRY(quant_stage(ctx, WS_IN2, q, n_q, &d_q));
  void *d_out = nullptr, *d_cnt = nullptr;
  if (n_q) LSG_TRY(lsg_ws(ctx, WS_OUT0, n_segments * n_q * 8 + 8, &d_out));
  if (count_out) LSG_TRY(lsg_ws(ctx, WS_OUT1, n_segments * 8 + 8, &d_cnt));
  LSG_TRY(lsg_segment_quantiles_dev(ctx, d_val, d_seg, n, n_segments, d_q, n_q, (double *)d_out, (uint64_t *)d_cnt));
  if (n_segments && n_q) LSG_CUDA(ctx, cudaMemcpyAsync(out, d_out, n_segments * n_q * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (n_segments && count_out) LSG_CUDA(ctx, cudaMemcpyAsync(count_out, d_cnt, n_segments * 8, cudaMemcpyDeviceToHost, ctx->stream));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

}  // extern "C"
