// synth_common.h — the synthetic workload of SURVEY.md §8(d): record i is a pure function of
// (seed, i), so the device generator (synth.cu) and the host generator (lsg_synth_host) produce the
// same bytes and any sub-range can be regenerated without materialising the whole input.
//
// Only integer arithmetic and IEEE multiplies/divides by small integers are used, so host and device
// agree bit for bit.  Value profiles follow pkg/signals/generator.go:244-289 (profileForFault), fault
// scenarios pkg/faultreplay/generator.go:10-16, multi-fault combinations pkg/benchmark/harness.go:228-241.
#pragma once
#include <stdint.h>

#include "../../include/llmslo_gpu.h"

#ifdef __CUDACC__
#define SYN_HD __host__ __device__ inline
#else
#define SYN_HD inline
#endif

SYN_HD uint64_t syn_mix(uint64_t x) {  // splitmix64 finaliser
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}
SYN_HD uint64_t syn_hash(uint64_t seed, uint64_t stream, uint64_t i) {
  return syn_mix(syn_mix(seed ^ (stream * 0xd6e8feb86659fd93ull)) ^ i);
}

// fault labels: 0 provider_throttle, 1 dns_latency, 2 cpu_throttle, 3 memory_pressure, 4 network_partition
// (pkg/faultreplay/generator.go:10-16) -> domain index in AllDomains order (mapper.go:29-50)
SYN_HD int syn_fault_domain(int f) {
  return f == 0 ? 4 : f == 1 ? 0 : f == 2 ? 2 : f == 3 ? 3 : 1;
}
SYN_HD int syn_domain_fault(int d) {   // inverse, for the multi-fault combinations (domains -> injected fault)
  return d == 4 ? 0 : d == 0 ? 1 : d == 2 ? 2 : d == 3 ? 3 : 4;
}

// profileForFault (signals/generator.go:244-289) for signal index s; fault -1 = baseline
SYN_HD double syn_profile(int fault, int s) {
  double base;
  switch (s) {
    case 0: base = 12; break;    // dns_latency_ms
    case 1: base = 0.2; break;   // tcp_retransmits_total
    case 2: base = 4; break;     // runqueue_delay_ms
    case 3: base = 18; break;    // connect_latency_ms
    case 4: base = 22; break;    // tls_handshake_ms
    case 5: base = 0.6; break;   // cpu_steal_pct
    case 6: base = 5; break;     // cfs_throttled_ms
    case 7: base = 0.5; break;   // mem_reclaim_latency_ms
    case 8: base = 2; break;     // disk_io_latency_ms
    case 9: base = 5; break;     // syscall_latency_ms
    default: base = 0; break;    // connect_errors_total, tls_handshake_fail_total
  }
  switch (fault) {
    case 1: if (s == 0) base = 220; if (s == 3) base = 130; break;
    case 2: if (s == 2) base = 28; if (s == 5) base = 9; if (s == 6) base = 170; break;
    case 3: if (s == 2) base = 14; if (s == 6) base = 90; if (s == 7) base = 25; if (s == 8) base = 60; break;
    case 0: if (s == 3) base = 45; if (s == 4) base = 55; if (s == 10) base = 1; if (s == 9) base = 250; break;
    case 4: if (s == 3) base = 350; if (s == 10) base = 3; if (s == 1) base = 12; if (s == 0) base = 180; if (s == 11) base = 2; break;
    default: break;
  }
  return base;
}

// multi-fault combinations of pkg/benchmark/harness.go:228-241 as (n, d0, d1, d2, is_mixed_label)
SYN_HD void syn_combo(int c, int *n, int dom[3], int *mixed) {
  const int tab[12][5] = {{2, 0, 2, 0, 1}, {2, 4, 3, 0, 1}, {2, 1, 4, 0, 1}, {2, 2, 3, 0, 1}, {2, 0, 3, 0, 1}, {3, 0, 4, 2, 1},
                          {1, 0, 0, 0, 0}, {1, 2, 0, 0, 0}, {1, 4, 0, 0, 0}, {1, 3, 0, 0, 0}, {1, 1, 0, 0, 0}, {2, 0, 1, 0, 1}};
  *n = tab[c][0];
  dom[0] = tab[c][1]; dom[1] = tab[c][2]; dom[2] = tab[c][3];
  *mixed = tab[c][4];
}

struct SynIncident {
  int n_faults;
  int fault[3];
  uint8_t actual;       // BuildConfusionMatrix's "actual": ExpectedDomain or MapFaultLabel(FaultLabel) (pipeline.go:73-77)
  uint32_t expected;    // ExpectedDomains as nibbles
  uint8_t expected_n;
};
SYN_HD SynIncident syn_incident(const lsg_synth_cfg &c, uint64_t k) {
  SynIncident r;
  const uint64_t gk = k + (uint64_t)c.rank * 0x100000000ull;   // distinct incidents per rank
  const uint64_t h = syn_hash(c.seed, 20, gk);
  if (!c.multi_fault) {
    const int f = (int)(h % 5);
    r.n_faults = 1; r.fault[0] = f; r.fault[1] = r.fault[2] = -1;
    r.actual = (uint8_t)syn_fault_domain(f);
    r.expected = r.actual; r.expected_n = 1;
  } else {
    int n, dom[3], mixed;
    syn_combo((int)(h % 12), &n, dom, &mixed);
    r.n_faults = n;
    r.expected = 0;
    for (int i = 0; i < 3; ++i) {
      r.fault[i] = i < n ? syn_domain_fault(dom[i]) : -1;
      if (i < n) r.expected |= (uint32_t)dom[i] << (4 * i);
    }
    r.expected_n = (uint8_t)n;
    r.actual = mixed ? 7 : (uint8_t)dom[0];   // fault_label "mixed" without expected_domain maps to "unknown"
  }
  return r;
}

SYN_HD lsg_ident syn_ident(const lsg_synth_cfg &c, uint32_t row) {
  lsg_ident r;
  const uint64_t gp = (uint64_t)c.rank * c.n_pods + row;      // global pod number
  const uint32_t per_node = c.pods_per_node / c.pods_per_group;  // services per node
  const uint64_t gg = gp / c.pods_per_group;                    // global (node, service) group
  r.pod = (uint32_t)(gp + 1);
  r.node = (uint32_t)(gp / c.pods_per_node + 1);
  r.service = (uint32_t)(gg % per_node + 1);
  r.grp = row / c.pods_per_group + 1;                           // rank-local numbering keeps the key short
  return r;
}

SYN_HD lsg_span syn_span(const lsg_synth_cfg &c, uint64_t i) {
  lsg_span s;
  const uint64_t h1 = syn_hash(c.seed, 1, i + (uint64_t)c.rank * 0x100000000ull);
  const uint64_t h2 = syn_hash(c.seed, 2, i + (uint64_t)c.rank * 0x100000000ull);
  s.ident = (uint32_t)(h1 % c.n_pods);
  s.ts_ns = c.t0_ns + (int64_t)(h2 % (uint64_t)c.horizon_ns);
  s.pid = 1000 + (uint32_t)((h1 >> 32) % 4);
  s.conn = 1 + (uint32_t)((h1 >> 40) % 8);
  s.trace = (uint32_t)(i + 1);
  s.incident = (uint32_t)(i / c.spans_per_incident);
  s.reserved = 0;
  return s;
}

SYN_HD lsg_event syn_event(const lsg_synth_cfg &c, uint64_t j) {
  const uint64_t gj = j + (uint64_t)c.rank * 0x100000000ull;
  const uint64_t hp = syn_hash(c.seed, 10, gj), hk = syn_hash(c.seed, 11, gj), ht = syn_hash(c.seed, 12, gj);
  const uint64_t hv = syn_hash(c.seed, 13, gj);
  const uint64_t parent = hp % c.n_spans;
  const lsg_span sp = syn_span(c, parent);
  lsg_event e;
  // which identity the event shares with its parent span
  const unsigned kind = (unsigned)(hk % 1000);
  e.ident = sp.ident; e.pid = sp.pid; e.conn = sp.conn;
  uint32_t trace = sp.trace;
  if (kind >= 500 && kind < 800) { trace = 0; e.conn = sp.conn + 100; }                 // pod + pid only
  else if (kind >= 800 && kind < 950) { trace = 0; e.pid = sp.pid + 100; }              // pod + conn only
  else if (kind >= 950) {                                                               // service + node only
    trace = 0; e.pid = sp.pid + 100; e.conn = sp.conn + 100;
    if (c.pods_per_group > 1) {
      const uint32_t g0 = sp.ident / c.pods_per_group * c.pods_per_group;
      e.ident = g0 + (sp.ident - g0 + 1) % c.pods_per_group;                            // the next replica of the group
    }
  }
  // |dt| mixture: 60% <=100 ms, 20% <=250 ms, 15% <=500 ms, 5% <=3 s; 1% exact tier edges +-1 ns; 2% on a 1 ms grid (ties)
  const unsigned m = (unsigned)(ht % 1000);
  const uint64_t lim = m < 600 ? 100000000ull : m < 800 ? 250000000ull : m < 950 ? 500000000ull : 3000000000ull;
  int64_t dt = (int64_t)((ht >> 10) % (2 * lim + 1)) - (int64_t)lim;
  const unsigned special = (unsigned)((hk >> 20) % 100);
  if (special == 0) {
    const int64_t edges[4] = {100000000ll, 250000000ll, 500000000ll, 2000000000ll};
    dt = edges[(hk >> 30) % 4] + (int64_t)((hk >> 34) % 3) - 1;
    if ((hk >> 40) & 1) dt = -dt;
  } else if (special < 3) {
    dt = dt / 1000000 * 1000000;
  }
  e.ts_ns = sp.ts_ns + dt;
  // signal type: the nine kernel probes (ebpf/c/llm_slo_event.h:5-15); 0.1% of an unsupported type
  const int sig9[9] = {0, 1, 2, 3, 4, 5, 7, 8, 9};
  unsigned sig = (unsigned)sig9[(hv >> 8) % 9];
  if ((hv >> 20) % 1000 == 0) sig = LSG_SIG_UNSUPPORTED;
  e.trace_sig = (trace << 4) | sig;
  // value: max of the injected faults' profiles for this signal, times a factor in [0.75, 1.25] on a 1/64 grid
  const SynIncident inc = syn_incident(c, sp.incident);
  double base = syn_profile(-1, (int)(sig & 15) < LSG_NSIG ? (int)sig : 0);
  for (int f = 0; f < inc.n_faults; ++f) {
    const double v = syn_profile(inc.fault[f], (int)(sig & 15) < LSG_NSIG ? (int)sig : 0);
    base = v > base ? v : base;
  }
  e.value = base * (double)(48 + (hv >> 32) % 33) / 64.0;
  return e;
}
