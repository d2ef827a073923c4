// synth_host.cpp — the synthetic workload generator (csrc/synth_common.h: record i is a pure function of (seed, i))
// compiled for the host alone, inside liboracle.so.  TEST INFRASTRUCTURE: it lets bench.py's `--impl reference`
// arm and the sampled-group checks build their inputs without mapping the product library.  Input production
// only, no path arithmetic; the same header generates the device copy, so the bytes are identical.
#include <thread>
#include <vector>

#include "../llm-slo-ebpf-toolkit_b200/csrc/synth_common.h"

extern "C" {

size_t orc_synth_n_incidents(const lsg_synth_cfg *c) {
  if (!c || c->spans_per_incident == 0) return 0;
  return (size_t)((c->n_spans + c->spans_per_incident - 1) / c->spans_per_incident);
}

// any of the output pointers may be NULL; events [ev_lo, ev_hi) and spans [sp_lo, sp_hi) of the configuration
int orc_synth_range(const lsg_synth_cfg *c, uint64_t sp_lo, uint64_t sp_hi, lsg_span *spans, uint64_t ev_lo, uint64_t ev_hi,
                    lsg_event *events, lsg_ident *idents, uint8_t *inc_actual, uint32_t *inc_expected, uint8_t *inc_expected_n,
                    int nthreads) {
  if (!c || c->n_pods == 0 || c->pods_per_node == 0 || c->pods_per_group == 0 || c->spans_per_incident == 0) return -1;
  if (c->pods_per_node % c->pods_per_group || c->n_pods % c->pods_per_group || c->horizon_ns <= 0 || c->n_spans == 0) return -1;
  if (nthreads < 1) nthreads = 1;
  auto par = [&](uint64_t lo, uint64_t hi, auto fn) {
    std::vector<std::thread> th;
    const uint64_t n = hi > lo ? hi - lo : 0, per = (n + nthreads - 1) / nthreads;
    for (int t = 0; t < nthreads; ++t) {
      const uint64_t a = lo + (uint64_t)t * per, b = a + per < hi ? a + per : hi;
      if (a < b) th.emplace_back([=]() { for (uint64_t i = a; i < b; ++i) fn(i); });
    }
    for (auto &x : th) x.join();
  };
  if (spans) par(sp_lo, sp_hi, [=](uint64_t i) { spans[i - sp_lo] = syn_span(*c, i); });
  if (events) par(ev_lo, ev_hi, [=](uint64_t j) { events[j - ev_lo] = syn_event(*c, j); });
  if (idents) for (uint32_t r = 0; r < c->n_pods; ++r) idents[r] = syn_ident(*c, r);
  const size_t nk = orc_synth_n_incidents(c);
  if (inc_actual || inc_expected || inc_expected_n)
    for (size_t k = 0; k < nk; ++k) {
      const SynIncident r = syn_incident(*c, k);
      if (inc_actual) inc_actual[k] = r.actual;
      if (inc_expected) inc_expected[k] = r.expected;
      if (inc_expected_n) inc_expected_n[k] = r.expected_n;
    }
  return 0;
}

}  // extern "C"
