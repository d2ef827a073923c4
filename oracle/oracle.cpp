// oracle.cpp — packed-record CPU restatement of the reference's correlation + attribution path.
//
// TEST INFRASTRUCTURE ONLY.  This is the checker the CUDA path is compared against; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.  The
// product (llm-slo-ebpf-toolkit_b200/, libllmslo_gpu.so) never links or calls it and has no CPU
// fallback.
//
// It works on the interned 32-byte records of include/llmslo_gpu.h.  The chain that pins it:
//   reference fixtures + *_test.go vectors  ->  oracle/ref_str.py (string-keyed, line-faithful)
//   ->  interner  ->  this file (tests/test_oracle_packed.py proves equality, fixtures + randomized)
//   ->  GPU (tests -m gpu).
// Parity status: PINNED (see oracle/ref_str.py header).  The reference cannot be compiled here (Go;
// no toolchain in the image), so there is no oracle/_ref.
//
// Paths in comments are relative to the reference root.
#include "../include/llmslo_gpu.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace {

const double kConf[5] = {0.0, 1.0, 0.9, 0.8, 0.65};  // dns.go:59,63,67,72
const int64_t kMs = 1000000;

// |a - b| as an exact unsigned magnitude; equals |float64(a.Sub(b))| of dns.go:111 whenever the
// window is < 2^53 ns (larger differences, including Sub's saturation, all compare false).
inline uint64_t absdiff(int64_t a, int64_t b) { return a > b ? (uint64_t)a - (uint64_t)b : (uint64_t)b - (uint64_t)a; }

// dns.go:107-113
inline bool within(int64_t a, int64_t b, int64_t window) {
  if (a == LSG_TS_ZERO || b == LSG_TS_ZERO) return false;
  return absdiff(a, b) <= (uint64_t)window;
}

// dns.go:50-76 on interned records.  Returns the tier (0 = Decision{}).
inline int match_tier(const lsg_span &s, const lsg_event &e, const lsg_ident *idents, int64_t window) {
  if (window <= 0) window = 2000 * kMs;
  if (!within(s.ts_ns, e.ts_ns, window)) return LSG_TIER_NONE;
  const uint32_t etrace = e.trace_sig >> 4;
  if (s.trace != 0 && s.trace == etrace) return LSG_TIER_TRACE;
  const lsg_ident &si = idents[s.ident];
  const lsg_ident &ei = idents[e.ident];
  if (si.pod != 0 && si.pod == ei.pod && s.pid > 0 && s.pid == e.pid && within(s.ts_ns, e.ts_ns, 100 * kMs))
    return LSG_TIER_POD_PID;
  if (si.pod != 0 && si.pod == ei.pod && s.conn != 0 && s.conn == e.conn && within(s.ts_ns, e.ts_ns, 250 * kMs))
    return LSG_TIER_POD_CONN;
  if (si.service != 0 && si.service == ei.service && si.node != 0 && si.node == ei.node &&
      within(s.ts_ns, e.ts_ns, 500 * kMs))
    return LSG_TIER_SVC_NODE;
  return LSG_TIER_NONE;
}

struct Cand {
  double conf;
  double dt;      // |span.ts - sig.ts| as float64, correlator.go:96-97
  uint32_t idx;   // input position of the signal
  uint8_t tier;
};

template <class F>
void parallel_for(size_t n, int nthreads, F f) {
  if (nthreads <= 1 || n < 2) {
    f((size_t)0, n, 0);
    return;
  }
  std::vector<std::thread> th;
  size_t chunk = (n + nthreads - 1) / nthreads;
  for (int t = 0; t < nthreads; ++t) {
    size_t lo = std::min(n, (size_t)t * chunk), hi = std::min(n, lo + chunk);
    th.emplace_back([=] { f(lo, hi, t); });
  }
  for (auto &x : th) x.join();
}

// correlator.go:92-125 given the candidates of one span in input order.
void finish_span(std::vector<Cand> &cands, const lsg_event *events, int fanout, const double *base, uint32_t *cand_event,
                 uint8_t *cand_tier, double *attr_out, double *conf_out, double *retrieval_out, uint32_t *n_cand_out,
                 uint64_t &fanout_dropped) {
  std::stable_sort(cands.begin(), cands.end(), [](const Cand &a, const Cand &b) {
    if (a.conf != b.conf) return a.conf > b.conf;
    return a.dt < b.dt;
  });
  if (n_cand_out) *n_cand_out = (uint32_t)cands.size();
  size_t keep = cands.size();
  if (keep > (size_t)fanout) {
    fanout_dropped += keep - fanout;
    keep = fanout;
  }
  const double nan = std::numeric_limits<double>::quiet_NaN();
  for (int s = 0; s < LSG_NSIG; ++s) attr_out[s] = base ? base[s] : nan;
  double maxc = 0.0;
  for (int k = 0; k < fanout; ++k) {
    cand_event[k] = LSG_NONE;
    cand_tier[k] = 0;
  }
  for (size_t k = 0; k < keep; ++k) {
    const lsg_event &e = events[cands[k].idx];
    int sig = e.trace_sig & 15;
    // `existing, ok := out[attr]; !ok || value > existing` (correlator.go:109)
    if (std::isnan(attr_out[sig]) || e.value > attr_out[sig]) attr_out[sig] = e.value;
    if (cands[k].conf > maxc) maxc = cands[k].conf;
    cand_event[k] = cands[k].idx;
    cand_tier[k] = cands[k].tier;
  }
  *conf_out = maxc;  // attribute is set only when > 0 (correlator.go:116-118)
  if (retrieval_out) {
    // DecomposeRetrieval, correlator.go:179-194: dns(0) + connect(3) + tls(4), in that order
    double total = 0.0;
    const int order[3] = {0, 3, 4};
    for (int k = 0; k < 3; ++k)
      if (!std::isnan(attr_out[order[k]])) total += attr_out[order[k]];
    *retrieval_out = total;
  }
}

}  // namespace

extern "C" {

int orc_match(const lsg_span *s, const lsg_event *e, const lsg_ident *idents, int64_t window_ns) {
  return match_tier(*s, *e, idents, window_ns);
}

// evaluator.go:90-167 (integer part) over pre-paired records
void orc_match_pairs(const lsg_span *spans, const lsg_event *events, size_t n, const lsg_ident *idents,
                     int64_t window_ns, double threshold, const uint8_t *expected_match, const uint8_t *expected_tier,
                     uint8_t *tier_out, uint8_t *predicted_out, lsg_eval *ev) {
  if (window_ns <= 0) window_ns = 2000 * kMs;
  if (threshold <= 0) threshold = 0.7;
  lsg_eval r;
  std::memset(&r, 0, sizeof r);
  for (size_t i = 0; i < n; ++i) {
    int tier = match_tier(spans[i], events[i], idents, window_ns);
    bool predicted = tier != 0 && kConf[tier] >= threshold;
    tier_out[i] = (uint8_t)tier;
    predicted_out[i] = predicted;
    r.matched_by_tier[tier]++;
    if (predicted) r.predicted_by_tier[tier]++;
    bool expm = expected_match ? expected_match[i] != 0 : false;
    if (expm && predicted) r.tp++;
    else if (!expm && predicted) r.fp++;
    else if (expm && !predicted) r.fn++;
    else r.tn++;
    uint8_t et = expected_tier ? expected_tier[i] : 0;
    if (expm && et != 0 && predicted) {
      r.tier_comparable++;
      if (et == tier) r.tier_correct++;
    }
  }
  *ev = r;
}

// processor.go:29-49 + correlator.go:50-125, literally: every span against every signal.
void orc_process_batch_brute(const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                             const lsg_ident *idents, int64_t window_ns, double threshold, int fanout,
                             const double *attr_base, uint32_t *cand_event, uint8_t *cand_tier, double *attr_out,
                             double *conf_out, double *retrieval_out, uint32_t *n_cand_out, lsg_debug *debug_out,
                             int nthreads) {
  if (window_ns <= 0) window_ns = 2000 * kMs;
  if (threshold <= 0) threshold = 0.7;
  if (fanout <= 0) fanout = 3;
  if (nthreads < 1) nthreads = 1;
  std::vector<lsg_debug> dbg(nthreads);
  for (auto &d : dbg) std::memset(&d, 0, sizeof d);
  parallel_for(n_spans, nthreads, [&](size_t lo, size_t hi, int t) {
    std::vector<Cand> cands;
    lsg_debug d;
    std::memset(&d, 0, sizeof d);
    for (size_t i = lo; i < hi; ++i) {
      cands.clear();
      const lsg_span &s = spans[i];
      for (size_t j = 0; j < n_events; ++j) {
        const lsg_event &e = events[j];
        if ((e.trace_sig & 15) >= LSG_NSIG) {  // signalAttrKey default branch
          d.unsupported_type++;
          continue;
        }
        int tier = match_tier(s, e, idents, window_ns);
        if (tier == 0) {
          d.unmatched++;
          continue;
        }
        if (kConf[tier] < threshold) {
          d.low_confidence++;
          continue;
        }
        cands.push_back(Cand{kConf[tier], (double)absdiff(s.ts_ns, e.ts_ns), (uint32_t)j, (uint8_t)tier});
      }
      finish_span(cands, events, fanout, attr_base ? attr_base + i * LSG_NSIG : nullptr, cand_event + i * fanout,
                  cand_tier + i * fanout, attr_out + i * LSG_NSIG, conf_out + i,
                  retrieval_out ? retrieval_out + i : nullptr, n_cand_out ? n_cand_out + i : nullptr, d.fanout_dropped);
    }
    dbg[t] = d;
  });
  lsg_debug total;
  std::memset(&total, 0, sizeof total);
  for (auto &d : dbg) {
    total.unmatched += d.unmatched;
    total.low_confidence += d.low_confidence;
    total.fanout_dropped += d.fanout_dropped;
    total.unsupported_type += d.unsupported_type;
  }
  *debug_out = total;
}

// Same results as orc_process_batch_brute, computed with an index: events ordered by (grp, ts), each
// span visits only its group's events inside the outer window.  Proven equal to the brute force in
// tests/test_oracle_packed.py; this is the "fair CPU" baseline (BASELINE.md §3 line B2).
void orc_process_batch_indexed(const lsg_span *spans, size_t n_spans, const lsg_event *events, size_t n_events,
                               const lsg_ident *idents, int64_t window_ns, double threshold, int fanout,
                               const double *attr_base, uint32_t *cand_event, uint8_t *cand_tier, double *attr_out,
                               double *conf_out, double *retrieval_out, uint32_t *n_cand_out, lsg_debug *debug_out,
                               int nthreads) {
  if (window_ns <= 0) window_ns = 2000 * kMs;
  if (threshold <= 0) threshold = 0.7;
  if (fanout <= 0) fanout = 3;
  if (nthreads < 1) nthreads = 1;
  // index of supported, non-zero-time events in groups != 0
  struct Ent { uint32_t grp; int64_t ts; uint32_t idx; };
  std::vector<Ent> ent;
  ent.reserve(n_events);
  uint64_t n_unsupported = 0, n_supported = 0;
  for (size_t j = 0; j < n_events; ++j) {
    const lsg_event &e = events[j];
    if ((e.trace_sig & 15) >= LSG_NSIG) { n_unsupported++; continue; }
    n_supported++;
    uint32_t g = idents[e.ident].grp;
    if (g == 0 || e.ts_ns == LSG_TS_ZERO) continue;
    ent.push_back(Ent{g, e.ts_ns, (uint32_t)j});
  }
  std::sort(ent.begin(), ent.end(), [](const Ent &a, const Ent &b) {
    if (a.grp != b.grp) return a.grp < b.grp;
    if (a.ts != b.ts) return a.ts < b.ts;
    return a.idx < b.idx;
  });
  std::vector<lsg_debug> dbg(nthreads);
  std::vector<uint64_t> matched(nthreads, 0);
  parallel_for(n_spans, nthreads, [&](size_t lo, size_t hi, int t) {
    std::vector<Cand> cands;
    lsg_debug d;
    std::memset(&d, 0, sizeof d);
    uint64_t nm = 0;
    for (size_t i = lo; i < hi; ++i) {
      cands.clear();
      const lsg_span &s = spans[i];
      uint32_t g = idents[s.ident].grp;
      if (g != 0 && s.ts_ns != LSG_TS_ZERO) {
        // saturating window bounds
        int64_t tlo = s.ts_ns < INT64_MIN + window_ns ? INT64_MIN : s.ts_ns - window_ns;
        int64_t thi = s.ts_ns > INT64_MAX - window_ns ? INT64_MAX : s.ts_ns + window_ns;
        auto first = std::lower_bound(ent.begin(), ent.end(), Ent{g, tlo, 0}, [](const Ent &a, const Ent &b) {
          if (a.grp != b.grp) return a.grp < b.grp;
          return a.ts < b.ts;
        });
        for (auto it = first; it != ent.end() && it->grp == g && it->ts <= thi; ++it) {
          const lsg_event &e = events[it->idx];
          int tier = match_tier(s, e, idents, window_ns);
          if (tier == 0) continue;
          nm++;
          if (kConf[tier] < threshold) { d.low_confidence++; continue; }
          cands.push_back(Cand{kConf[tier], (double)absdiff(s.ts_ns, e.ts_ns), it->idx, (uint8_t)tier});
        }
        // restore input order so that the stable sort of finish_span breaks ties as the reference does
        std::sort(cands.begin(), cands.end(), [](const Cand &a, const Cand &b) { return a.idx < b.idx; });
      }
      finish_span(cands, events, fanout, attr_base ? attr_base + i * LSG_NSIG : nullptr, cand_event + i * fanout,
                  cand_tier + i * fanout, attr_out + i * LSG_NSIG, conf_out + i,
                  retrieval_out ? retrieval_out + i : nullptr, n_cand_out ? n_cand_out + i : nullptr, d.fanout_dropped);
    }
    dbg[t] = d;
    matched[t] = nm;
  });
  lsg_debug total;
  std::memset(&total, 0, sizeof total);
  uint64_t nm = 0;
  for (int t = 0; t < nthreads; ++t) {
    total.low_confidence += dbg[t].low_confidence;
    total.fanout_dropped += dbg[t].fanout_dropped;
    nm += matched[t];
  }
  total.unsupported_type = n_unsupported * (uint64_t)n_spans;
  total.unmatched = n_supported * (uint64_t)n_spans - nm;
  *debug_out = total;
}

// A7 glue (no reference counterpart beyond "max", correlator.go:109-111): per incident and signal,
// over the incident's spans in span order: max, count, sum, and a log2 histogram of the values
// (bin = clamp(floor(log2(v)) + 4, 0, 15) for v > 0; v <= 0 -> bin 0).
void orc_incident_features(const double *span_attr, const uint32_t *span_incident, size_t n_spans, size_t n_incidents,
                           double *signals_out, lsg_feature *features_out) {
  const double nan = std::numeric_limits<double>::quiet_NaN();
  for (size_t k = 0; k < n_incidents * LSG_NSIG; ++k) signals_out[k] = nan;
  if (features_out) std::memset(features_out, 0, sizeof(lsg_feature) * n_incidents * LSG_NSIG);
  for (size_t i = 0; i < n_spans; ++i) {
    uint32_t inc = span_incident[i];
    if (inc == LSG_NONE || inc >= n_incidents) continue;
    for (int s = 0; s < LSG_NSIG; ++s) {
      double v = span_attr[i * LSG_NSIG + s];
      if (std::isnan(v)) continue;
      double &m = signals_out[(size_t)inc * LSG_NSIG + s];
      if (std::isnan(m) || v > m) m = v;
      if (features_out) {
        lsg_feature &f = features_out[(size_t)inc * LSG_NSIG + s];
        f.count++;
        f.sum += v;
        int bin = 0;
        if (v > 0) {
          int ex;
          std::frexp(v, &ex);  // v = m * 2^ex, m in [0.5, 1)  ->  floor(log2 v) = ex - 1
          bin = ex - 1 + 4;
          bin = bin < 0 ? 0 : (bin > LSG_SKETCH_BINS - 1 ? LSG_SKETCH_BINS - 1 : bin);
        }
        f.sketch[bin]++;
      }
    }
  }
}

void orc_default_model(lsg_model *m) {
  // bayesian.go:54-207
  static const double lik[LSG_NSIG][LSG_NDOM] = {
      {0.95, 0.70, 0.10, 0.10, 0.10, 0.10, 0.15, 0.10}, {0.15, 0.90, 0.10, 0.10, 0.10, 0.15, 0.10, 0.10},
      {0.10, 0.10, 0.90, 0.60, 0.10, 0.10, 0.10, 0.10}, {0.50, 0.85, 0.10, 0.10, 0.75, 0.40, 0.30, 0.10},
      {0.10, 0.30, 0.10, 0.10, 0.80, 0.50, 0.20, 0.10}, {0.10, 0.10, 0.90, 0.20, 0.10, 0.10, 0.10, 0.10},
      {0.10, 0.10, 0.85, 0.75, 0.10, 0.10, 0.10, 0.10}, {0.05, 0.05, 0.15, 0.95, 0.05, 0.05, 0.05, 0.05},
      {0.05, 0.05, 0.10, 0.85, 0.05, 0.05, 0.30, 0.05}, {0.10, 0.20, 0.15, 0.10, 0.90, 0.60, 0.40, 0.10},
      {0.10, 0.80, 0.05, 0.05, 0.60, 0.85, 0.15, 0.10}, {0.05, 0.70, 0.05, 0.05, 0.30, 0.60, 0.10, 0.05}};
  static const double thr[LSG_NSIG] = {40, 2, 10, 80, 60, 2, 40, 5, 10, 50, 1, 1};
  std::memset(m, 0, sizeof *m);
  for (int d = 0; d < LSG_NDOM; ++d) m->prior[d] = 1.0 / 8.0;
  std::memcpy(m->lik, lik, sizeof lik);
  std::memcpy(m->thresh, thr, sizeof thr);
  m->sig_mask = 0xFFF;
}

static inline double clampl(double p) { return p < 0.01 ? 0.01 : (p > 0.99 ? 0.99 : p); }  // bayesian.go:303-311

// bayesian.go:218-284 + the numeric part of AttributeSample (bayesian.go:315-343).
// Fixed summation order: log prior, signals 0..11 (those in sig_mask), log_extra; LSE in AllDomains order.
void orc_attribute(const lsg_model *m, const double *signals, const uint8_t *has_signals, const uint8_t *rule_label,
                   size_t n, double *posterior, uint8_t *top1, uint16_t *elev_mask) {
  const double nan = std::numeric_limits<double>::quiet_NaN();
  for (size_t i = 0; i < n; ++i) {
    const double *sg = signals + i * LSG_NSIG;
    bool has;
    if (has_signals) has = has_signals[i] != 0;
    else {
      has = false;
      for (int s = 0; s < LSG_NSIG; ++s) has |= !std::isnan(sg[s]);
    }
    if (!has) {
      for (int d = 0; d < LSG_NDOM; ++d) posterior[i * LSG_NDOM + d] = nan;
      top1[i] = rule_label ? rule_label[i] : 0xFF;
      if (elev_mask) elev_mask[i] = 0;
      continue;
    }
    unsigned elev = 0;
    for (int s = 0; s < LSG_NSIG; ++s)
      if (!std::isnan(sg[s]) && !std::isnan(m->thresh[s]) && sg[s] >= m->thresh[s]) elev |= 1u << s;
    double logp[LSG_NDOM];
    for (int d = 0; d < LSG_NDOM; ++d) {
      double prior = m->prior[d];
      if (!(prior > 0)) prior = 1e-10;
      double lp = std::log(prior);
      for (int s = 0; s < LSG_NSIG; ++s) {
        if (!(m->sig_mask >> s & 1)) continue;
        double p = m->lik[s][d];
        double l = std::isnan(p) ? 0.5 : ((elev >> s & 1) ? clampl(p) : clampl(1 - p));
        lp += std::log(l);
      }
      lp += m->log_extra[d];
      logp[d] = lp;
    }
    double mx = -INFINITY;
    for (int d = 0; d < LSG_NDOM; ++d)
      if (logp[d] > mx) mx = logp[d];
    double se = 0.0;
    for (int d = 0; d < LSG_NDOM; ++d) se += std::exp(logp[d] - mx);
    double logz = mx + std::log(se);
    int best = 0;
    double bestp = -1.0;
    for (int d = 0; d < LSG_NDOM; ++d) {
      double p = std::exp(logp[d] - logz);
      posterior[i * LSG_NDOM + d] = p;
      if (p > bestp) { bestp = p; best = d; }  // first maximum = sort.SliceStable order
    }
    top1[i] = (uint8_t)best;
    if (elev_mask) elev_mask[i] = (uint16_t)elev;
  }
}

// pipeline.go:65-185 (integer part)
void orc_confusion(const uint8_t *actual, const uint8_t *predicted, const uint32_t *expected, const uint8_t *expected_n,
                   const double *posterior, double coverage_thr, double hyp_floor, size_t n, lsg_confusion_out *out) {
  std::memset(out, 0, sizeof *out);
  for (size_t i = 0; i < n; ++i) {
    uint8_t a = actual[i], p = predicted[i];
    if (a < LSG_MAX_ACTUAL && p < LSG_NDOM) out->matrix[a][p]++;
    out->n++;
    if (a == p) out->correct++;
    unsigned hyp = 0;
    if (posterior)
      for (int d = 0; d < LSG_NDOM; ++d) {
        double q = posterior[i * LSG_NDOM + d];
        if (!std::isnan(q) && !(q < hyp_floor) && q >= coverage_thr) hyp |= 1u << d;  // bayesian.go:326-328, pipeline.go:163-167
      }
    if (p < LSG_NDOM) hyp |= 1u << p;
    int en = expected_n[i], found = 0;
    bool part = false;
    for (int k = 0; k < en; ++k) {
      unsigned dcode = expected[i] >> (4 * k) & 15;
      if (dcode < LSG_NDOM) {
        if (dcode == p) part = true;
        if (hyp >> dcode & 1) found++;
      }
    }
    if (part) out->partial++;
    out->coverage[found][en]++;
  }
}

// retry_storm.go:46-62 replayed over a batch: events are (pod, ts) in arrival order; out[i] = count after
// Record(pod_i, ts_i) (prefix-only pruning, retry_storm.go:98-110) and storm[i] = count >= threshold.
void orc_retry_storm(const uint32_t *pod, const int64_t *ts, size_t n, uint32_t n_pods, int64_t window_ns, int threshold,
                     uint32_t *count_out, uint8_t *storm_out) {
  std::vector<std::vector<int64_t>> buckets(n_pods);
  std::vector<size_t> start(n_pods, 0);
  for (size_t i = 0; i < n; ++i) {
    auto &b = buckets[pod[i]];
    size_t &st = start[pod[i]];
    b.push_back(ts[i]);
    int64_t cutoff = ts[i] - window_ns;
    while (st < b.size() && b[st] < cutoff) st++;
    uint32_t c = (uint32_t)(b.size() - st);
    count_out[i] = c;
    storm_out[i] = c >= (uint32_t)threshold;
  }
}

// quantile(), pkg/slo/calculator.go:121-149 (same body pkg/releasegate/gate.go:737-760), per segment.
// sort.Float64s orders NaN before every number; ids >= n_segments are skipped.
void orc_segment_quantiles(const double *values, const uint32_t *segment, size_t n, size_t n_segments, const double *q,
                           size_t n_q, double *out, uint64_t *count_out) {
  std::vector<std::vector<double>> lists(n_segments);
  for (size_t i = 0; i < n; ++i) {
    const uint32_t s = segment ? segment[i] : 0u;
    if (s < n_segments) lists[s].push_back(values[i]);
  }
  for (size_t s = 0; s < n_segments; ++s) {
    auto &sorted = lists[s];
    std::stable_sort(sorted.begin(), sorted.end(), [](double a, double b) { return a < b || (a != a && b == b); });
    if (count_out) count_out[s] = sorted.size();
    for (size_t j = 0; j < n_q; ++j) {
      double qq = q[j], r;
      if (sorted.empty()) {
        r = 0;
      } else {
        if (qq <= 0) qq = 0;
        if (qq >= 1) qq = 1;
        if (sorted.size() == 1) {
          r = sorted[0];
        } else {
          const double pos = qq * (double)(sorted.size() - 1);
          const long lower = (long)std::floor(pos), upper = (long)std::ceil(pos);
          if (lower == upper) {
            r = sorted[lower];
          } else {
            const double frac = pos - (double)lower;
            r = sorted[lower] * (1 - frac) + sorted[upper] * frac;
          }
        }
      }
      out[s * n_q + j] = r;
    }
  }
}

int orc_hw_threads(void) { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
