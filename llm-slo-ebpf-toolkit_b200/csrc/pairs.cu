// pairs.cu — K-pairs: correlation.Match (pkg/correlation/dns.go:50-76) over n pre-paired records and the
// integer part of EvaluateLabeledPairs (pkg/correlation/evaluator.go:115-152).
//
// HBM-bound streaming kernel: 64 B in + 2 B out per pair, one thread per pair, 128-bit loads, counters by
// warp ballot -> shared memory -> one global atomic per counter per block.
#include "common.cuh"

namespace {

constexpr int kEvalWords = sizeof(lsg_eval) / 8;  // 16

__global__ void __launch_bounds__(256)
k_match_pairs(const lsg_span *__restrict__ spans, const lsg_event *__restrict__ events, size_t n,
              const lsg_ident *__restrict__ idents, uint32_t n_idents, int64_t window_ns, double threshold,
              const uint8_t *__restrict__ expected_match, const uint8_t *__restrict__ expected_tier,
              uint8_t *__restrict__ tier_out, uint8_t *__restrict__ predicted_out, unsigned long long *__restrict__ eval,
              unsigned int *__restrict__ bad) {
  __shared__ unsigned int s_cnt[kEvalWords];
  if (threadIdx.x < kEvalWords) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31;
  for (size_t base = (size_t)blockIdx.x * blockDim.x; base < n; base += (size_t)gridDim.x * blockDim.x) {
    size_t i = base + threadIdx.x;
    bool valid = i < n;
    int tier = 0;
    bool predicted = false, expm = false;
    unsigned et = 0;
    if (valid) {
      SpanRec s = load_span(spans + i);
      EventRec e = load_event(events + i);
      if (s.ident >= n_idents || e.ident >= n_idents) {
        atomicOr(bad, 1u);
      } else if (s.ts != LSG_TS_ZERO && e.ts != LSG_TS_ZERO) {   // withinWindow, dns.go:107-113
        uint64_t adt = lsg_absdiff(s.ts, e.ts);
        if (adt <= (uint64_t)window_ns) {
          lsg_ident si = load_ident(idents, s.ident), ei = load_ident(idents, e.ident);
          tier = lsg_match_tier(adt, s.trace, s.pid, s.conn, si, e.trace_sig >> 4, e.pid, e.conn, ei);
        }
      }
      predicted = tier != 0 && lsg_tier_conf(tier) >= threshold;   // evaluator.go:117
      expm = expected_match ? expected_match[i] != 0 : false;
      et = expected_tier ? expected_tier[i] : 0;
      tier_out[i] = (uint8_t)tier;
      predicted_out[i] = predicted ? 1 : 0;
    }
    // evaluator.go:134-150, one ballot per counter
    const unsigned full = 0xffffffffu;
    unsigned b_tp = __ballot_sync(full, valid && expm && predicted);
    unsigned b_fp = __ballot_sync(full, valid && !expm && predicted);
    unsigned b_fn = __ballot_sync(full, valid && expm && !predicted);
    unsigned b_tn = __ballot_sync(full, valid && !expm && !predicted);
    bool comparable = valid && expm && et != 0 && predicted;
    unsigned b_cmp = __ballot_sync(full, comparable);
    unsigned b_cor = __ballot_sync(full, comparable && et == (unsigned)tier);
    unsigned b_pt[5], b_mt[5];
#pragma unroll
    for (int t = 0; t < 5; ++t) {
      b_pt[t] = __ballot_sync(full, valid && predicted && tier == t);
      b_mt[t] = __ballot_sync(full, valid && tier == t);
    }
    if (lane == 0) {
      if (b_tp) atomicAdd(&s_cnt[0], __popc(b_tp));
      if (b_fp) atomicAdd(&s_cnt[1], __popc(b_fp));
      if (b_fn) atomicAdd(&s_cnt[2], __popc(b_fn));
      if (b_tn) atomicAdd(&s_cnt[3], __popc(b_tn));
      if (b_cor) atomicAdd(&s_cnt[4], __popc(b_cor));
      if (b_cmp) atomicAdd(&s_cnt[5], __popc(b_cmp));
#pragma unroll
      for (int t = 0; t < 5; ++t) {
        if (b_pt[t]) atomicAdd(&s_cnt[6 + t], __popc(b_pt[t]));
        if (b_mt[t]) atomicAdd(&s_cnt[11 + t], __popc(b_mt[t]));
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < kEvalWords && s_cnt[threadIdx.x]) atomicAdd(&eval[threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]);
}

}  // namespace

int lsg_launch_match_pairs(lsg_ctx *ctx, const lsg_span *spans, const lsg_event *events, size_t n, const lsg_ident *idents,
                           size_t n_idents, int64_t window_ns, double threshold, const uint8_t *expected_match,
                           const uint8_t *expected_tier, uint8_t *tier_out, uint8_t *predicted_out, uint64_t *eval_dev) {
  // eval_dev: kEvalWords counters followed by one u32 "bad ident" flag word, zeroed here
  LSG_CUDA(ctx, cudaMemsetAsync(eval_dev, 0, (kEvalWords + 1) * 8, ctx->stream));
  if (n == 0) return 0;
  size_t blocks = (n + 255) / 256;
  size_t cap = (size_t)ctx->sm_count * 16;   // grid-stride: a multiple of the SM count, <= 16 CTAs/SM of work queued
  if (blocks > cap) blocks = cap;
  ProfScope ps(ctx, PK_PAIRS);
  k_match_pairs<<<(unsigned)blocks, 256, 0, ctx->stream>>>(spans, events, n, idents, (uint32_t)n_idents, window_ns, threshold,
                                                           expected_match, expected_tier, tier_out, predicted_out,
                                                           (unsigned long long *)eval_dev,
                                                           (unsigned int *)(eval_dev + kEvalWords));
  LSG_LAUNCH_CHECK(ctx);
  return 0;
}
