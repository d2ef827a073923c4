// validate.cu — lsg_validate_groups: a check of the JOIN-GROUP CONTRACT of include/llmslo_gpu.h on a caller's batch.
//
// The kernels only pair records that carry the same non-zero join group, so a caller whose groups are too fine gets
// silently fewer matches than the reference's Match() (pkg/correlation/dns.go:50-76) would give.  Match() can accept
// a pair on three identities: the trace id (dns.go:58-60), the pod (dns.go:61-68) and the (service, node) pair
// (dns.go:69-73).  The contract therefore holds exactly when each of those identities appears under ONE group:
//   - identity rows that share a non-empty pod carry the same group,
//   - identity rows that share a non-empty (service, node) pair carry the same group,
//   - records (spans and supported, timed events) that share a non-zero trace id carry the same group.
// Three open-addressing tables (first group seen per key, atomicCAS); a key met again under another group is a conflict.
#include "common.cuh"

namespace {

constexpr uint32_t VG_EMPTY = 0xffffffffu;

__device__ __forceinline__ uint32_t vg_mix(uint64_t k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return (uint32_t)k;
}

// first group seen for `key` (non-zero) in a table of `mask + 1` slots; returns true when `grp` disagrees with it
__device__ __forceinline__ bool vg_conflict(unsigned long long *keys, uint32_t *grps, uint32_t mask, unsigned long long key, uint32_t grp) {
  uint32_t slot = vg_mix(key) & mask;
  for (;;) {
    const unsigned long long prev = atomicCAS(&keys[slot], 0ull, key);
    if (prev == 0ull || prev == key) break;
    slot = (slot + 1) & mask;
  }
  const uint32_t old = atomicCAS(&grps[slot], VG_EMPTY, grp);
  return old != VG_EMPTY && old != grp;
}

__global__ void __launch_bounds__(256)
k_vg_idents(const lsg_ident *__restrict__ idents, uint32_t n, unsigned long long *pod_keys, uint32_t *pod_grps,
            unsigned long long *cls_keys, uint32_t *cls_grps, uint32_t mask, unsigned long long *out) {
  for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
    const lsg_ident id = load_ident(idents, r);
    if (id.pod != 0 && vg_conflict(pod_keys, pod_grps, mask, (unsigned long long)id.pod, id.grp)) atomicAdd(&out[0], 1ull);
    if (id.service != 0 && id.node != 0 &&
        vg_conflict(cls_keys, cls_grps, mask, ((unsigned long long)id.service << 32) | id.node, id.grp))
      atomicAdd(&out[1], 1ull);
  }
}

template <bool SPANS>
__global__ void __launch_bounds__(256)
k_vg_traces(const lsg_event *__restrict__ recs, size_t n, const lsg_ident *__restrict__ idents, uint32_t n_idents,
            unsigned long long *keys, uint32_t *grps, uint32_t mask, unsigned long long *out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const U8 u = ld_nc_v8(recs + i);
    const int64_t ts = (int64_t)(((uint64_t)u.a.y << 32) | u.a.x);
    uint32_t ident, trace;
    if (SPANS) {
      ident = u.a.z; trace = u.b.y;
    } else {
      ident = u.b.x; trace = u.b.w >> 4;
      if ((u.b.w & 15u) >= LSG_NSIG) continue;   // an unsupported signal never reaches Match() (correlator.go:73-77)
    }
    if (ident >= n_idents) { atomicAdd(&out[3], 1ull); continue; }
    if (trace == 0 || ts == LSG_TS_ZERO) continue;   // no trace id / the zero Time: cannot match on this identity
    if (vg_conflict(keys, grps, mask, (unsigned long long)trace, load_ident(idents, ident).grp)) atomicAdd(&out[2], 1ull);
  }
}

}  // namespace

extern "C" {

int lsg_validate_groups_dev(lsg_ctx *ctx, const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                            const lsg_ident *idents, size_t n_idents, uint64_t out[4]) {
  if (!ctx || !out) return LSG_ERR_INVALID;
  if ((n_spans && !spans) || (n_events && !events) || ((n_spans || n_events || n_idents) && !idents))
    return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  if (n_idents >= ((size_t)1 << 31) || n_spans + n_events >= ((size_t)1 << 31))
    return lsg_fail(ctx, LSG_ERR_RANGE, "batch too large for the validation tables");
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  size_t icap = 64, tcap = 64;
  while (icap < 2 * n_idents) icap <<= 1;
  while (tcap < 2 * (n_spans + n_events)) tcap <<= 1;
  void *iv, *tv, *small_v;
  LSG_TRY(lsg_ws(ctx, WS_IDENT_TABLE, 2 * icap * 12, &iv));     // (the join's class table is rebuilt on its next call)
  LSG_TRY(lsg_ws(ctx, WS_KEYS_A, tcap * 12, &tv));
  LSG_TRY(lsg_ws(ctx, WS_SMALL, 4096, &small_v));
  unsigned long long *pod_keys = (unsigned long long *)iv, *cls_keys = pod_keys + icap;
  uint32_t *pod_grps = (uint32_t *)(cls_keys + icap), *cls_grps = pod_grps + icap;
  unsigned long long *tr_keys = (unsigned long long *)tv;
  uint32_t *tr_grps = (uint32_t *)(tr_keys + tcap);
  unsigned long long *out_dev = (unsigned long long *)((char *)small_v + 2048);
  LSG_CUDA(ctx, cudaMemsetAsync(pod_keys, 0, 2 * icap * 8, ctx->stream));
  LSG_CUDA(ctx, cudaMemsetAsync(pod_grps, 0xff, 2 * icap * 4, ctx->stream));
  LSG_CUDA(ctx, cudaMemsetAsync(tr_keys, 0, tcap * 8, ctx->stream));
  LSG_CUDA(ctx, cudaMemsetAsync(tr_grps, 0xff, tcap * 4, ctx->stream));
  LSG_CUDA(ctx, cudaMemsetAsync(out_dev, 0, 4 * 8, ctx->stream));
  auto grid = [&](size_t n) { size_t b = (n + 255) / 256, cap = (size_t)ctx->sm_count * 8; return (unsigned)(b > cap ? cap : (b ? b : 1)); };
  if (n_idents) {
    k_vg_idents<<<grid(n_idents), 256, 0, ctx->stream>>>(idents, (uint32_t)n_idents, pod_keys, pod_grps, cls_keys, cls_grps,
                                                          (uint32_t)icap - 1, out_dev);
    LSG_LAUNCH_CHECK(ctx);
  }
  if (n_spans) {
    k_vg_traces<true><<<grid(n_spans), 256, 0, ctx->stream>>>((const lsg_event *)spans, n_spans, idents, (uint32_t)n_idents, tr_keys,
                                                               tr_grps, (uint32_t)tcap - 1, out_dev);
    LSG_LAUNCH_CHECK(ctx);
  }
  if (n_events) {
    k_vg_traces<false><<<grid(n_events), 256, 0, ctx->stream>>>(events, n_events, idents, (uint32_t)n_idents, tr_keys, tr_grps,
                                                                 (uint32_t)tcap - 1, out_dev);
    LSG_LAUNCH_CHECK(ctx);
  }
  LSG_CUDA(ctx, cudaMemcpyAsync(ctx->pinned, out_dev, 4 * 8, cudaMemcpyDeviceToHost, ctx->stream));
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(out, ctx->pinned, 4 * 8);
  return 0;
}

int lsg_validate_groups(lsg_ctx *ctx, const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                        const lsg_ident *idents, size_t n_idents, uint64_t out[4]) {
  if (!ctx || !out) return LSG_ERR_INVALID;
  if ((n_spans && !spans) || (n_events && !events) || ((n_spans || n_events || n_idents) && !idents))
    return lsg_fail(ctx, LSG_ERR_INVALID, "null argument");
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  void *ds = nullptr, *de = nullptr, *di = nullptr;
  LSG_TRY(lsg_ws(ctx, WS_SPANS_IN, n_spans * sizeof(lsg_span), &ds));
  LSG_TRY(lsg_ws(ctx, WS_EVENTS_IN, n_events * sizeof(lsg_event), &de));
  LSG_TRY(lsg_ws(ctx, WS_IDENTS, n_idents * sizeof(lsg_ident), &di));
  if (n_spans) LSG_CUDA(ctx, cudaMemcpyAsync(ds, spans, n_spans * sizeof(lsg_span), cudaMemcpyHostToDevice, ctx->stream));
  if (n_events) LSG_CUDA(ctx, cudaMemcpyAsync(de, events, n_events * sizeof(lsg_event), cudaMemcpyHostToDevice, ctx->stream));
  if (n_idents) LSG_CUDA(ctx, cudaMemcpyAsync(di, idents, n_idents * sizeof(lsg_ident), cudaMemcpyHostToDevice, ctx->stream));
  return lsg_validate_groups_dev(ctx, (const lsg_span *)ds, n_spans, (const lsg_event *)de, n_events, (const lsg_ident *)di, n_idents, out);
}

}  // extern "C"
