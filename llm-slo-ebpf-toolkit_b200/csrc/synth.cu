// synth.cu — device and host generators of the synthetic workload (see synth_common.h).
#include "common.cuh"
#include "synth_common.h"

namespace {
__global__ void __launch_bounds__(256) k_synth_spans(const __grid_constant__ lsg_synth_cfg c, lsg_span *out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < c.n_spans; i += (size_t)gridDim.x * blockDim.x) {
    lsg_span s = syn_span(c, i);
    uint4 a, b;
    a.x = (uint32_t)(uint64_t)s.ts_ns; a.y = (uint32_t)((uint64_t)s.ts_ns >> 32); a.z = s.ident; a.w = s.pid;
    b.x = s.conn; b.y = s.trace; b.z = s.incident; b.w = 0;
    st_v8(out + i, U8{a, b});
  }
}
__global__ void __launch_bounds__(256) k_synth_events(const __grid_constant__ lsg_synth_cfg c, lsg_event *out) {
  for (size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x; j < c.n_events; j += (size_t)gridDim.x * blockDim.x) {
    lsg_event e = syn_event(c, j);
    uint4 a, b;
    const uint64_t vb = (uint64_t)__double_as_longlong(e.value);
    a.x = (uint32_t)(uint64_t)e.ts_ns; a.y = (uint32_t)((uint64_t)e.ts_ns >> 32); a.z = (uint32_t)vb; a.w = (uint32_t)(vb >> 32);
    b.x = e.ident; b.y = e.pid; b.z = e.conn; b.w = e.trace_sig;
    st_v8(out + j, U8{a, b});
  }
}
__global__ void __launch_bounds__(256)
k_synth_misc(const __grid_constant__ lsg_synth_cfg c, lsg_ident *idents, size_t n_idents, uint8_t *inc_actual,
             uint32_t *inc_expected, uint8_t *inc_expected_n, size_t n_inc) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_idents + n_inc; i += (size_t)gridDim.x * blockDim.x) {
    if (i < n_idents) {
      if (idents) idents[i] = syn_ident(c, (uint32_t)i);
    } else {
      const size_t k = i - n_idents;
      const SynIncident r = syn_incident(c, k);
      if (inc_actual) inc_actual[k] = r.actual;
      if (inc_expected) inc_expected[k] = r.expected;
      if (inc_expected_n) inc_expected_n[k] = r.expected_n;
    }
  }
}
int check_cfg(const lsg_synth_cfg *c) {
  if (!c || c->n_pods == 0 || c->pods_per_node == 0 || c->pods_per_group == 0 || c->spans_per_incident == 0) return -1;
  if (c->pods_per_node % c->pods_per_group || c->n_pods % c->pods_per_group) return -1;
  if (c->horizon_ns <= 0 || c->n_spans == 0) return -1;
  if (c->n_spans >= (1ull << LSG_TRACE_BITS)) return -1;   // one trace id per span
  return 0;
}
}  // namespace

extern "C" size_t lsg_synth_n_idents(const lsg_synth_cfg *cfg) { return cfg ? cfg->n_pods : 0; }
extern "C" size_t lsg_synth_n_incidents(const lsg_synth_cfg *cfg) {
  if (!cfg || cfg->spans_per_incident == 0) return 0;
  return (size_t)((cfg->n_spans + cfg->spans_per_incident - 1) / cfg->spans_per_incident);
}

extern "C" int lsg_synth_dev(lsg_ctx *ctx, const lsg_synth_cfg *cfg, lsg_span *spans, lsg_event *events, lsg_ident *idents,
                             uint8_t *inc_actual, uint32_t *inc_expected, uint8_t *inc_expected_n) {
  if (!ctx) return LSG_ERR_INVALID;
  if (check_cfg(cfg)) return lsg_fail(ctx, LSG_ERR_INVALID, "invalid synthetic workload configuration");
  LSG_CUDA(ctx, cudaSetDevice(ctx->device));
  const unsigned cap = (unsigned)ctx->sm_count * 8;
  auto grid = [&](size_t n) { size_t b = (n + 255) / 256; return (unsigned)(b > cap ? cap : (b ? b : 1)); };
  if (spans) { k_synth_spans<<<grid(cfg->n_spans), 256, 0, ctx->stream>>>(*cfg, spans); LSG_LAUNCH_CHECK(ctx); }
  if (events && cfg->n_events) { k_synth_events<<<grid(cfg->n_events), 256, 0, ctx->stream>>>(*cfg, events); LSG_LAUNCH_CHECK(ctx); }
  const size_t ni = lsg_synth_n_idents(cfg), nk = lsg_synth_n_incidents(cfg);
  k_synth_misc<<<grid(ni + nk), 256, 0, ctx->stream>>>(*cfg, idents, ni, inc_actual, inc_expected, inc_expected_n, nk);
  LSG_LAUNCH_CHECK(ctx);
  LSG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int lsg_synth_host(const lsg_synth_cfg *cfg, lsg_span *spans, lsg_event *events, lsg_ident *idents,
                              uint8_t *inc_actual, uint32_t *inc_expected, uint8_t *inc_expected_n) {
  if (check_cfg(cfg)) return LSG_ERR_INVALID;
  if (spans) for (uint64_t i = 0; i < cfg->n_spans; ++i) spans[i] = syn_span(*cfg, i);
  if (events) for (uint64_t j = 0; j < cfg->n_events; ++j) events[j] = syn_event(*cfg, j);
  if (idents) for (uint32_t r = 0; r < cfg->n_pods; ++r) idents[r] = syn_ident(*cfg, r);
  const size_t nk = lsg_synth_n_incidents(cfg);
  for (size_t k = 0; k < nk; ++k) {
    const SynIncident r = syn_incident(*cfg, k);
    if (inc_actual) inc_actual[k] = r.actual;
    if (inc_expected) inc_expected[k] = r.expected;
    if (inc_expected_n) inc_expected_n[k] = r.expected_n;
  }
  return 0;
}
