// llmslo.hpp — C++ host layer over the C ABI (include/llmslo_gpu.h), mirroring the reference's Go packages for the
// hot path: same type names, field names, function names, argument meaning, defaults and empty-input behaviour.
//
//   llmslo::correlation      pkg/correlation            (dns.go, evaluator.go)
//   llmslo::ebpfcorrelator   pkg/otel/processor/ebpfcorrelator (correlator.go, processor.go)
//   llmslo::attribution      pkg/attribution            (bayesian.go, pipeline.go, mapper.go)
//
// The reference is Go; no Go toolchain exists where this was built, so the compiled host side is C++ (the cgo
// binding a maintainer would write is in INTEGRATION.md).  Everything that computes runs in libllmslo_gpu.so; this
// header only interns strings on the way in and rebuilds maps/strings on the way out.  No CPU fallback: without a
// GPU the first call throws.
#pragma once
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <map>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../../include/llmslo_gpu.h"

namespace llmslo {

// time.Time / time.Duration: Unix nanoseconds; the zero Time is its own value (dns.go:108 IsZero)
struct Time {
  int64_t ns = LSG_TS_ZERO;
  Time() = default;
  explicit Time(int64_t unix_ns) : ns(unix_ns) {}
  bool IsZero() const { return ns == LSG_TS_ZERO; }
  Time Add(int64_t d) const { return IsZero() ? *this : Time(ns + d); }
};
using Duration = int64_t;
constexpr Duration Millisecond = 1000000, Second = 1000000000;

// one lsg_ctx per process for the free functions below
class Context {
 public:
  static lsg_ctx *get() {
    static Context c;
    return c.h_;
  }
  static void check(int rc) {
    if (rc != 0) throw std::runtime_error(std::string("libllmslo_gpu: ") + lsg_last_error(get()));
  }

 private:
  Context() {
    if (lsg_create(0, nullptr, &h_) != 0) throw std::runtime_error("lsg_create failed: no usable CUDA device (there is no CPU path)");
  }
  ~Context() { lsg_destroy(h_); }
  lsg_ctx *h_ = nullptr;
};

namespace detail {

inline const std::vector<std::string> &signals() {   // DefaultLikelihoods order, bayesian.go:69-189
  static const std::vector<std::string> v = {"dns_latency_ms", "tcp_retransmits_total", "runqueue_delay_ms", "connect_latency_ms",
                                             "tls_handshake_ms", "cpu_steal_pct", "cfs_throttled_ms", "mem_reclaim_latency_ms",
                                             "disk_io_latency_ms", "syscall_latency_ms", "connect_errors_total",
                                             "tls_handshake_fail_total"};
  return v;
}
inline const std::vector<std::string> &attr_keys() {   // signalAttrKey, correlator.go:143-172 (pkg/semconv/llm_ebpf.go)
  static const std::vector<std::string> v = {"llm.ebpf.dns.latency_ms", "llm.ebpf.tcp.retransmits", "llm.ebpf.sched.runqueue_delay_ms",
                                             "llm.ebpf.net.connect_latency_ms", "llm.ebpf.tls.handshake_ms", "llm.ebpf.cpu.steal_pct",
                                             "llm.ebpf.cpu.cfs_throttled_ms", "llm.ebpf.mm.reclaim_latency_ms", "llm.ebpf.blk.io_latency_ms",
                                             "llm.ebpf.syscall.latency_ms", "llm.ebpf.net.connect_errors_total",
                                             "llm.ebpf.tls.handshake_fail_total"};
  return v;
}
inline const std::vector<std::string> &domains() {   // AllDomains, bayesian.go:23-34
  static const std::vector<std::string> v = {"network_dns", "network_egress", "cpu_throttle", "memory_pressure", "provider_throttle",
                                             "provider_error", "retrieval_backend", "unknown"};
  return v;
}
inline int index_of(const std::vector<std::string> &v, const std::string &s) {
  for (size_t i = 0; i < v.size(); ++i)
    if (v[i] == s) return (int)i;
  return -1;
}
inline const char *tier_name(int t) {
  static const char *n[5] = {"", "trace_id_exact", "pod_pid_100ms", "pod_conn_250ms", "service_node_500ms"};
  return n[t];
}
inline double tier_conf(int t) {
  static const double c[5] = {0.0, 1.0, 0.9, 0.8, 0.65};
  return c[t];
}

// Dictionary encoding + the join groups the header's contract asks for (union-find over identity rows)
class Interner {
 public:
  template <class Rec>
  uint32_t ident(const Rec &r) {
    const uint32_t pod = id(pod_, r.Pod), node = id(node_, r.Node), svc = id(svc_, r.Service);
    // a batch has few distinct identities and consecutive records often repeat the previous one: the row of the last
    // triple is kept, everything else is integer-keyed (no strings are built per record)
    if (have_last_ && pod == last_.pod && node == last_.node && svc == last_.service) return last_row_;
    const Triple key{pod, node, svc};
    auto it = rows_.find(key);
    uint32_t row;
    if (it != rows_.end()) {
      row = it->second;
    } else {
      row = (uint32_t)idents_.size();
      rows_.emplace(key, row);
      idents_.push_back(lsg_ident{pod, node, svc, 0});
      parent_.push_back(row);
      if (pod) link(by_pod_, (uint64_t)pod, row);
      if (node && svc) link(by_sn_, ((uint64_t)svc << 32) | node, row);
    }
    last_ = lsg_ident{pod, node, svc, 0}; last_row_ = row; have_last_ = true;
    return row;
  }
  uint32_t trace(const std::string &t, uint32_t row) {
    const uint32_t tid = id(trace_, t);
    if (tid >= (1u << LSG_TRACE_BITS)) throw std::runtime_error("more than 2^28 distinct trace ids in one batch");
    if (tid) {   // first row seen for a trace id, then union with every later one
      if (tid >= trace_row_.size()) trace_row_.resize((size_t)tid + 1024, 0xffffffffu);
      uint32_t &first = trace_row_[tid];
      if (first == 0xffffffffu) first = row;
      else if (first != row) unite(first, row);
    }
    return tid;
  }
  uint32_t conn(const std::string &c) { return id(conn_, c); }
  static uint32_t pid(int64_t p) { return p <= 0 ? 0u : (uint32_t)p; }
  std::vector<lsg_ident> table() {
    std::unordered_map<uint32_t, uint32_t> comp;
    std::vector<lsg_ident> out = idents_;
    for (uint32_t r = 0; r < out.size(); ++r) {
      const uint32_t root = find(r);
      auto it = comp.find(root);
      if (it == comp.end()) it = comp.emplace(root, (uint32_t)comp.size() + 1).first;
      out[r].grp = it->second;
    }
    if (out.empty()) out.push_back(lsg_ident{0, 0, 0, 0});
    return out;
  }

 private:
  using Dict = std::unordered_map<std::string, uint32_t>;
  static uint32_t id(Dict &d, const std::string &s) {
    if (s.empty()) return 0;
    auto it = d.find(s);
    if (it != d.end()) return it->second;
    const uint32_t v = (uint32_t)d.size() + 1;
    d[s] = v;
    return v;
  }
  uint32_t find(uint32_t x) {
    while (parent_[x] != x) x = parent_[x] = parent_[parent_[x]];
    return x;
  }
  void unite(uint32_t x, uint32_t y) {
    const uint32_t a = find(x), b = find(y);
    if (a != b) parent_[std::max(a, b)] = std::min(a, b);
  }
  void link(std::unordered_map<uint64_t, uint32_t> &d, uint64_t key, uint32_t row) {
    auto it = d.find(key);
    if (it == d.end()) { d.emplace(key, row); return; }
    unite(it->second, row);
  }
  struct Triple {
    uint32_t pod, node, service;
    bool operator==(const Triple &o) const { return pod == o.pod && node == o.node && service == o.service; }
  };
  struct TripleHash {
    size_t operator()(const Triple &t) const {
      uint64_t h = ((uint64_t)t.pod << 32 | t.node) * 0x9e3779b97f4a7c15ull;
      return (size_t)((h ^ (h >> 29)) + t.service * 0xbf58476d1ce4e5b9ull);
    }
  };
  Dict trace_, pod_, node_, svc_, conn_;
  std::unordered_map<Triple, uint32_t, TripleHash> rows_;
  std::unordered_map<uint64_t, uint32_t> by_pod_, by_sn_;
  std::vector<uint32_t> trace_row_;
  lsg_ident last_{0, 0, 0, 0};
  uint32_t last_row_ = 0;
  bool have_last_ = false;
  std::vector<lsg_ident> idents_;
  std::vector<uint32_t> parent_;
};

}  // namespace detail

// ------------------------------------------------------------------------------------------------------------
namespace correlation {

constexpr Duration DefaultWindow = 2 * Second;             // dns.go:12
constexpr double DefaultEnrichmentThreshold = 0.7;         // dns.go:15

struct SpanRef {   // dns.go:19-27
  std::string TraceID, Service, Node, Pod;
  int64_t PID = 0;
  std::string ConnTuple;
  Time Timestamp;
};
struct SignalRef {   // dns.go:30-40
  std::string Signal, TraceID, Service, Node, Pod;
  int64_t PID = 0;
  std::string ConnTuple;
  Time Timestamp;
  double Value = 0;
};
struct Decision {   // dns.go:43-47
  bool Matched = false;
  double Confidence = 0;
  std::string Tier;
};
struct LabeledPair {   // evaluator.go:13-19
  std::string CaseID;
  SpanRef Span;
  SignalRef Signal;
  bool ExpectedMatch = false;
  std::string ExpectedTier;
};
struct Prediction {   // evaluator.go:22-31
  std::string CaseID;
  bool Expected = false, Predicted = false;
  double Confidence = 0;
  std::string Tier;
  bool Correct = false;
  std::string Signal, ExpectedTier;
};
struct EvalReport {   // evaluator.go:34-51
  int SampleSize = 0, TruePositive = 0, FalsePositive = 0, FalseNegative = 0, TrueNegative = 0;
  double Precision = 0, Recall = 0, F1 = 0, TierAccuracy = 0, MeanConfidence = 0;
  int WindowMS = 0;
  double Threshold = 0;
};
struct GateResult { bool Pass; std::string Message; };   // evaluator.go:54-57

inline lsg_span pack_span(detail::Interner &in, const SpanRef &s) {
  const uint32_t row = in.ident(s);
  return lsg_span{s.Timestamp.ns, row, detail::Interner::pid(s.PID), in.conn(s.ConnTuple), in.trace(s.TraceID, row), LSG_NONE, 0};
}
inline lsg_event pack_event(detail::Interner &in, const SignalRef &g) {
  const uint32_t row = in.ident(g);
  const int sig = detail::index_of(detail::signals(), g.Signal);
  return lsg_event{g.Timestamp.ns, g.Value, row, detail::Interner::pid(g.PID), in.conn(g.ConnTuple),
                   (in.trace(g.TraceID, row) << 4) | (uint32_t)(sig < 0 ? LSG_SIG_UNSUPPORTED : sig)};
}
inline Decision decision_of(int tier) {
  if (tier == 0) return Decision{};
  return Decision{true, detail::tier_conf(tier), detail::tier_name(tier)};
}

// evaluator.go:90-167.  Per-pair decisions and the integer counters come from the GPU; the ratios are formed here.
inline std::pair<EvalReport, std::vector<Prediction>> EvaluateLabeledPairs(const std::vector<LabeledPair> &pairs, Duration window,
                                                                           double threshold) {
  if (window <= 0) window = DefaultWindow;
  if (threshold <= 0) threshold = DefaultEnrichmentThreshold;
  EvalReport report;
  report.SampleSize = (int)pairs.size();
  report.WindowMS = (int)(window / Millisecond);
  report.Threshold = threshold;
  std::vector<Prediction> out;
  if (pairs.empty()) return {report, out};
  detail::Interner in;
  std::vector<lsg_span> sp;
  std::vector<lsg_event> ev;
  std::vector<uint8_t> em, et, tier(pairs.size()), pred(pairs.size());
  for (auto &p : pairs) {
    sp.push_back(pack_span(in, p.Span));
    ev.push_back(pack_event(in, p.Signal));
    em.push_back(p.ExpectedMatch);
    int t = 0;
    if (!p.ExpectedTier.empty()) {
      t = 7;   // an unknown tier name can never equal decision.Tier
      for (int k = 1; k <= 4; ++k)
        if (p.ExpectedTier == detail::tier_name(k)) t = k;
    }
    et.push_back((uint8_t)t);
  }
  auto idents = in.table();
  lsg_eval e;
  Context::check(lsg_match_pairs(Context::get(), sp.data(), ev.data(), pairs.size(), idents.data(), idents.size(), window, threshold,
                                 em.data(), et.data(), tier.data(), pred.data(), &e));
  for (size_t i = 0; i < pairs.size(); ++i) {
    const Decision d = decision_of(tier[i]);
    out.push_back(Prediction{pairs[i].CaseID, pairs[i].ExpectedMatch, pred[i] != 0, d.Confidence, d.Tier,
                             (pred[i] != 0) == pairs[i].ExpectedMatch, pairs[i].Signal.Signal, pairs[i].ExpectedTier});
  }
  auto safe_div = [](uint64_t a, uint64_t b) { return b == 0 ? 0.0 : (double)a / (double)b; };   // evaluator.go:189-194
  report.TruePositive = (int)e.tp; report.FalsePositive = (int)e.fp; report.FalseNegative = (int)e.fn; report.TrueNegative = (int)e.tn;
  report.Precision = safe_div(e.tp, e.tp + e.fp);
  report.Recall = safe_div(e.tp, e.tp + e.fn);
  if (report.Precision + report.Recall > 0) report.F1 = 2 * ((report.Precision * report.Recall) / (report.Precision + report.Recall));
  if (e.tier_comparable > 0) report.TierAccuracy = (double)e.tier_correct / (double)e.tier_comparable;
  uint64_t n_pred = 0;
  double conf_sum = 0;
  for (int t = 0; t < 5; ++t) { n_pred += e.predicted_by_tier[t]; conf_sum += (double)e.predicted_by_tier[t] * detail::tier_conf(t); }
  if (n_pred > 0) report.MeanConfidence = conf_sum / (double)n_pred;
  return {report, out};
}

// dns.go:50-76
inline Decision Match(const SpanRef &span, const SignalRef &signal, Duration window) {
  auto r = EvaluateLabeledPairs({LabeledPair{"", span, signal, false, ""}}, window, 1e-300);
  return Decision{!r.second[0].Tier.empty(), r.second[0].Confidence, r.second[0].Tier};
}

// evaluator.go:170-187
inline GateResult EvaluateGate(const EvalReport &report, double minPrecision, double minRecall) {
  char buf[128];
  if (report.Precision < minPrecision) {
    snprintf(buf, sizeof buf, "precision gate failed: got %.4f required %.4f", report.Precision, minPrecision);
    return {false, buf};
  }
  if (report.Recall < minRecall) {
    snprintf(buf, sizeof buf, "recall gate failed: got %.4f required %.4f", report.Recall, minRecall);
    return {false, buf};
  }
  return {true, "correlation gate passed"};
}

}  // namespace correlation

// ------------------------------------------------------------------------------------------------------------
namespace ebpfcorrelator {

using AttrMap = std::map<std::string, double>;
const std::string AttrCorrelationConf = "llm.ebpf.correlation_confidence";
const std::string AttrRetrievalKernelMS = "llm.ebpf.retrieval.kernel_attributed_ms";

struct DebugStats { int64_t Unmatched = 0, LowConfidence = 0, FanoutDropped = 0, UnsupportedType = 0; };   // correlator.go:20-25
struct Candidate { correlation::SignalRef Signal; correlation::Decision Decision; int Index = -1; };      // correlator.go:28-31
struct EnrichmentResult { AttrMap Attributes; std::vector<Candidate> Candidates; DebugStats Debug; };       // correlator.go:34-38
struct SpanRecord {   // processor.go:10-20
  std::string TraceID, SpanID, Service, Node, Pod;
  int64_t PID = 0;
  std::string ConnTuple;
  Time Timestamp;
  AttrMap Attributes;
  std::vector<Candidate> Candidates;   // not in the Go struct: kept for diagnostics
};
struct ProcessedBatch { std::vector<SpanRecord> Spans; DebugStats Debug; };   // processor.go:23-26

// correlator.go:179-194 for a caller-owned map (inside ProcessBatch the sum comes from the GPU)
inline double DecomposeRetrieval(AttrMap &attrs) {
  double total = 0;
  for (int k : {0, 3, 4}) {
    auto it = attrs.find(detail::attr_keys()[k]);
    if (it != attrs.end()) total += it->second;
  }
  if (total > 0) attrs[AttrRetrievalKernelMS] = total;
  return total;
}

struct Correlator {   // correlator.go:13-17
  int WindowMS = 2000;
  double EnrichmentThreshold = 0.7;
  int MaxJoinFanout = 3;

  // processor.go:29-49: ONE C-ABI call for the whole batch
  ProcessedBatch ProcessBatch(const std::vector<SpanRecord> &spans, const std::vector<correlation::SignalRef> &signals) const {
    return run(spans, signals, true);
  }
  // correlator.go:50-125
  EnrichmentResult EnrichAttributes(const AttrMap &base, const correlation::SpanRef &span,
                                    const std::vector<correlation::SignalRef> &signals) const {
    SpanRecord r;
    r.TraceID = span.TraceID; r.Service = span.Service; r.Node = span.Node; r.Pod = span.Pod; r.PID = span.PID;
    r.ConnTuple = span.ConnTuple; r.Timestamp = span.Timestamp; r.Attributes = base;
    ProcessedBatch b = run({r}, signals, false);
    return EnrichmentResult{b.Spans[0].Attributes, b.Spans[0].Candidates, b.Debug};
  }

 private:
  ProcessedBatch run(const std::vector<SpanRecord> &spans, const std::vector<correlation::SignalRef> &signals, bool decompose) const {
    ProcessedBatch out;
    if (spans.empty()) return out;
    const auto t_begin = std::chrono::steady_clock::now();
    const int fanout = MaxJoinFanout > 0 ? MaxJoinFanout : 3;
    detail::Interner in;
    std::vector<lsg_span> sp;
    std::vector<lsg_event> ev;
    const double nan = std::nan("");
    std::vector<double> base(spans.size() * LSG_NSIG, nan);
    bool any_base = false;
    for (size_t i = 0; i < spans.size(); ++i) {
      const SpanRecord &s = spans[i];
      correlation::SpanRef ref{s.TraceID, s.Service, s.Node, s.Pod, s.PID, s.ConnTuple, s.Timestamp};
      sp.push_back(correlation::pack_span(in, ref));
      for (auto &kv : s.Attributes) {
        const int slot = detail::index_of(detail::attr_keys(), kv.first);
        if (slot >= 0) { base[i * LSG_NSIG + slot] = kv.second; any_base = true; }
      }
    }
    for (auto &g : signals) ev.push_back(correlation::pack_event(in, g));
    auto idents = in.table();
    std::vector<uint32_t> ce(spans.size() * fanout);
    std::vector<uint8_t> ct(spans.size() * fanout);
    std::vector<double> attr(spans.size() * LSG_NSIG), conf(spans.size()), retr(spans.size());
    lsg_debug dbg;
    const auto t_packed = std::chrono::steady_clock::now();
    Context::check(lsg_process_batch(Context::get(), sp.data(), sp.size(), ev.data(), ev.size(), idents.data(), idents.size(),
                                     (int64_t)WindowMS * Millisecond, EnrichmentThreshold, fanout, any_base ? base.data() : nullptr,
                                     ce.data(), ct.data(), attr.data(), conf.data(), retr.data(), nullptr, &dbg));
    const auto t_gpu = std::chrono::steady_clock::now();
    for (size_t i = 0; i < spans.size(); ++i) {
      SpanRecord r = spans[i];
      for (int k = 0; k < LSG_NSIG; ++k)
        if (!std::isnan(attr[i * LSG_NSIG + k])) r.Attributes[detail::attr_keys()[k]] = attr[i * LSG_NSIG + k];
      if (conf[i] > 0) r.Attributes[AttrCorrelationConf] = conf[i];
      if (decompose && retr[i] > 0) r.Attributes[AttrRetrievalKernelMS] = retr[i];
      r.Candidates.clear();
      for (int k = 0; k < fanout; ++k) {
        const uint32_t e = ce[i * fanout + k];
        if (e != LSG_NONE) r.Candidates.push_back(Candidate{signals[e], correlation::decision_of(ct[i * fanout + k]), (int)e});
      }
      out.Spans.push_back(std::move(r));
    }
    out.Debug = DebugStats{(int64_t)dbg.unmatched, (int64_t)dbg.low_confidence, (int64_t)dbg.fanout_dropped, (int64_t)dbg.unsupported_type};
    // where a call from the string signatures spends its time (read by host_tests --time)
    const auto t_end = std::chrono::steady_clock::now();
    auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    last_timing() = Timing{ms(t_begin, t_packed), ms(t_packed, t_gpu), ms(t_gpu, t_end)};
    return out;
  }

 public:
  struct Timing { double intern_ms, library_ms, rebuild_ms; };   // strings -> records | lsg_process_batch | outputs -> maps
  static Timing &last_timing() {
    static Timing t{0, 0, 0};
    return t;
  }
};
inline Correlator New() { return Correlator{}; }   // correlator.go:41-47

}  // namespace ebpfcorrelator

// ------------------------------------------------------------------------------------------------------------
namespace attribution {

using SignalMap = std::map<std::string, double>;

struct Posterior { std::string Domain; double Posterior; std::vector<std::string> Evidence; };   // bayesian.go:210-214
struct FaultSample {   // mapper.go:11-26 (the fields this path reads)
  std::string IncidentID, FaultLabel, ExpectedDomain;
  std::vector<std::string> ExpectedDomains;
  SignalMap Signals;
  double Confidence = 0;
};
struct FaultHypothesis { std::string Domain; double Posterior; std::vector<std::string> Evidence; };
struct IncidentAttribution {   // the fields of schema.IncidentAttribution this path writes
  std::string IncidentID, PredictedFaultDomain;
  double Confidence = 0;
  std::vector<FaultHypothesis> FaultHypotheses;
};
struct MatrixKey {   // pipeline.go:8-11
  std::string Actual, Predicted;
  bool operator<(const MatrixKey &o) const { return Actual != o.Actual ? Actual < o.Actual : Predicted < o.Predicted; }
};

inline std::vector<std::string> AllDomains() { return detail::domains(); }

// mapper.go:29-50
inline std::string MapFaultLabel(const std::string &label) {
  static const std::map<std::string, std::string> m = {
      {"dns_latency", "network_dns"}, {"egress_drop", "network_egress"}, {"cpu_throttle", "cpu_throttle"},
      {"memory_pressure", "memory_pressure"}, {"network_partition", "network_egress"}, {"provider_throttle", "provider_throttle"},
      {"provider_error", "provider_error"}, {"retrieval_slowdown", "retrieval_backend"}};
  auto it = m.find(label);
  return it == m.end() ? "unknown" : it->second;
}

struct BayesianAttributor {   // bayesian.go:39-52; Priors and Likelihoods are public and mutable as in Go
  std::map<std::string, double> Priors;
  std::map<std::string, std::map<std::string, double>> Likelihoods;

  BayesianAttributor() {
    lsg_model m;
    lsg_default_model(&m);
    for (int d = 0; d < LSG_NDOM; ++d) Priors[detail::domains()[d]] = m.prior[d];
    for (int s = 0; s < LSG_NSIG; ++s)
      for (int d = 0; d < LSG_NDOM; ++d) Likelihoods[detail::signals()[s]][detail::domains()[d]] = m.lik[s][d];
  }

  // bayesian.go:218-284
  std::vector<Posterior> Attribute(const SignalMap &signals) const {
    std::vector<double> post;
    std::vector<uint8_t> top1;
    std::vector<uint16_t> elev;
    run({signals}, {1}, {0}, post, top1, elev);
    return sorted_posteriors(post.data(), elev[0]);
  }
  // bayesian.go:315-343 for a batch: one launch for all incidents
  std::vector<IncidentAttribution> AttributeSamples(const std::vector<FaultSample> &samples) const {
    std::vector<IncidentAttribution> out;
    if (samples.empty()) return out;
    std::vector<SignalMap> maps;
    std::vector<uint8_t> has, rule;
    for (auto &s : samples) {
      maps.push_back(s.Signals);
      has.push_back(!s.Signals.empty());
      rule.push_back((uint8_t)detail::index_of(detail::domains(), MapFaultLabel(s.FaultLabel)));
    }
    std::vector<double> post;
    std::vector<uint8_t> top1;
    std::vector<uint16_t> elev;
    run(maps, has, rule, post, top1, elev);
    for (size_t i = 0; i < samples.size(); ++i) {
      IncidentAttribution a{samples[i].IncidentID, MapFaultLabel(samples[i].FaultLabel), samples[i].Confidence, {}};   // BuildAttribution
      if (!samples[i].Signals.empty()) {
        for (auto &p : sorted_posteriors(post.data() + i * LSG_NDOM, elev[i]))
          if (!(p.Posterior < 0.01)) a.FaultHypotheses.push_back(FaultHypothesis{p.Domain, p.Posterior, p.Evidence});
        a.PredictedFaultDomain = detail::domains()[top1[i]];
        a.Confidence = post[i * LSG_NDOM + top1[i]];
      }
      out.push_back(std::move(a));
    }
    return out;
  }
  IncidentAttribution AttributeSample(const FaultSample &s) const { return AttributeSamples({s})[0]; }

 private:
  lsg_model model() const {
    lsg_model m, def;
    lsg_default_model(&def);
    std::fill((double *)&m, (double *)&m + sizeof(m) / sizeof(double), 0.0);
    m.sig_mask = 0;
    m.reserved = 0;
    const double nan = std::nan("");
    for (int d = 0; d < LSG_NDOM; ++d) {
      auto it = Priors.find(detail::domains()[d]);
      m.prior[d] = it == Priors.end() ? 0.0 : it->second;
      m.log_extra[d] = 0.0;
    }
    for (int s = 0; s < LSG_NSIG; ++s) {
      m.thresh[s] = def.thresh[s];
      auto ll = Likelihoods.find(detail::signals()[s]);
      if (ll != Likelihoods.end()) m.sig_mask |= 1u << s;
      for (int d = 0; d < LSG_NDOM; ++d) {
        m.lik[s][d] = nan;
        if (ll != Likelihoods.end()) {
          auto it = ll->second.find(detail::domains()[d]);
          if (it != ll->second.end()) m.lik[s][d] = it->second;
        }
      }
    }
    for (auto &kv : Likelihoods) {   // rows outside the 12 known signals can never be elevated: a constant per domain
      if (detail::index_of(detail::signals(), kv.first) >= 0) continue;
      for (int d = 0; d < LSG_NDOM; ++d) {
        auto it = kv.second.find(detail::domains()[d]);
        double p = 0.5;
        if (it != kv.second.end()) p = std::min(std::max(1.0 - it->second, 0.01), 0.99);
        m.log_extra[d] += std::log(p);
      }
    }
    return m;
  }
  void run(const std::vector<SignalMap> &maps, const std::vector<uint8_t> &has, const std::vector<uint8_t> &rule,
           std::vector<double> &post, std::vector<uint8_t> &top1, std::vector<uint16_t> &elev) const {
    const size_t n = maps.size();
    const lsg_model m = model();
    Context::check(lsg_set_model(Context::get(), &m));
    std::vector<double> sig(n * LSG_NSIG, std::nan(""));
    for (size_t i = 0; i < n; ++i)
      for (auto &kv : maps[i]) {
        const int s = detail::index_of(detail::signals(), kv.first);
        if (s >= 0) sig[i * LSG_NSIG + s] = kv.second;
      }
    post.assign(n * LSG_NDOM, 0.0);
    top1.assign(n, 0);
    elev.assign(n, 0);
    Context::check(lsg_attribute(Context::get(), sig.data(), has.data(), rule.data(), n, post.data(), top1.data(), elev.data()));
  }
  std::vector<Posterior> sorted_posteriors(const double *post, uint16_t elev) const {
    std::vector<Posterior> out;
    for (int d = 0; d < LSG_NDOM; ++d) {
      std::vector<std::string> evid;
      for (int s = 0; s < LSG_NSIG; ++s) {
        if (!((elev >> s) & 1)) continue;
        auto ll = Likelihoods.find(detail::signals()[s]);
        if (ll == Likelihoods.end()) continue;
        auto it = ll->second.find(detail::domains()[d]);
        if (it != ll->second.end() && it->second >= 0.5) evid.push_back(detail::signals()[s]);
      }
      std::sort(evid.begin(), evid.end());
      out.push_back(Posterior{detail::domains()[d], post[d], evid});
    }
    std::stable_sort(out.begin(), out.end(), [](const Posterior &a, const Posterior &b) { return a.Posterior > b.Posterior; });   // bayesian.go:279-281
    return out;
  }
};
inline BayesianAttributor NewBayesianAttributor() { return BayesianAttributor(); }

// pipeline.go:20-62
inline std::vector<IncidentAttribution> BuildAttributionsRule(const std::vector<FaultSample> &samples) {
  std::vector<IncidentAttribution> out;
  for (auto &s : samples) out.push_back(IncidentAttribution{s.IncidentID, MapFaultLabel(s.FaultLabel), s.Confidence, {}});
  return out;
}
inline std::vector<IncidentAttribution> BuildAttributionsBayes(const std::vector<FaultSample> &samples, const BayesianAttributor *a = nullptr) {
  BayesianAttributor def;
  return (a ? *a : def).AttributeSamples(samples);
}
inline std::vector<IncidentAttribution> BuildAttributions(const std::vector<FaultSample> &samples, std::string mode) {
  std::string m;
  for (char c : mode)
    if (!isspace((unsigned char)c)) m.push_back((char)tolower((unsigned char)c));
  return m == "rule" ? BuildAttributionsRule(samples) : BuildAttributionsBayes(samples);
}

namespace detail2 {
struct Encoded {
  size_t n = 0;
  std::vector<uint8_t> actual, pred, exp_n;
  std::vector<uint32_t> exp;
  std::vector<double> post;
  std::vector<std::string> names;
};
inline Encoded encode(const std::vector<FaultSample> &samples, const std::vector<IncidentAttribution> &preds) {
  Encoded e;
  e.n = std::min(samples.size(), preds.size());
  e.names = detail::domains();
  e.names.resize(LSG_MAX_ACTUAL);
  std::map<std::string, int> other;
  e.post.assign(e.n * LSG_NDOM, std::nan(""));
  for (size_t i = 0; i < e.n; ++i) {
    const std::string actual = samples[i].ExpectedDomain.empty() ? MapFaultLabel(samples[i].FaultLabel) : samples[i].ExpectedDomain;
    int a = detail::index_of(detail::domains(), actual);
    if (a < 0) {
      auto it = other.find(actual);
      if (it == other.end()) {
        if (other.size() >= (size_t)(LSG_MAX_ACTUAL - LSG_NDOM)) throw std::runtime_error("too many actual classes outside AllDomains");
        it = other.emplace(actual, LSG_NDOM + (int)other.size()).first;
        e.names[it->second] = actual;
      }
      a = it->second;
    }
    e.actual.push_back((uint8_t)a);
    const int p = detail::index_of(detail::domains(), preds[i].PredictedFaultDomain);
    if (p < 0) throw std::runtime_error("predicted domain is not one of AllDomains");
    e.pred.push_back((uint8_t)p);
    std::vector<std::string> expected = samples[i].ExpectedDomains;
    if (expected.empty()) expected.push_back(actual);
    if (expected.size() > 8) throw std::runtime_error("more than 8 expected domains");
    uint32_t code = 0;
    for (size_t k = 0; k < expected.size(); ++k) {
      const int d = detail::index_of(detail::domains(), expected[k]);
      code |= (uint32_t)(d < 0 ? 0xF : d) << (4 * k);
    }
    e.exp.push_back(code);
    e.exp_n.push_back((uint8_t)expected.size());
    for (auto &h : preds[i].FaultHypotheses) e.post[i * LSG_NDOM + detail::index_of(detail::domains(), h.Domain)] = h.Posterior;
  }
  return e;
}
inline lsg_confusion_out confusion(const Encoded &e, double thr) {
  lsg_confusion_out out;
  Context::check(lsg_confusion(Context::get(), e.actual.data(), e.pred.data(), e.exp.data(), e.exp_n.data(), e.post.data(), thr, 0.0,
                               e.n, &out));
  return out;
}
}  // namespace detail2

// pipeline.go:65-82
inline std::map<MatrixKey, int> BuildConfusionMatrix(const std::vector<FaultSample> &samples, const std::vector<IncidentAttribution> &preds) {
  std::map<MatrixKey, int> m;
  if (samples.empty() || preds.empty()) return m;
  const auto e = detail2::encode(samples, preds);
  const auto out = detail2::confusion(e, 0.0);
  for (int a = 0; a < LSG_MAX_ACTUAL; ++a)
    for (int p = 0; p < LSG_NDOM; ++p)
      if (out.matrix[a][p]) m[MatrixKey{e.names[a], detail::domains()[p]}] = (int)out.matrix[a][p];
  return m;
}
// pipeline.go:85-104 (the denominator is len(predictions), as in Go)
inline double Accuracy(const std::vector<FaultSample> &samples, const std::vector<IncidentAttribution> &preds) {
  if (preds.empty() || samples.empty()) return 0;
  return (double)detail2::confusion(detail2::encode(samples, preds), 0.0).correct / (double)preds.size();
}
// pipeline.go:110-136
inline double PartialAccuracy(const std::vector<FaultSample> &samples, const std::vector<IncidentAttribution> &preds) {
  if (preds.empty() || samples.empty()) return 0;
  return (double)detail2::confusion(detail2::encode(samples, preds), 0.0).partial / (double)preds.size();
}
// pipeline.go:140-185
inline double CoverageAccuracy(const std::vector<FaultSample> &samples, const std::vector<IncidentAttribution> &preds, double threshold) {
  if (preds.empty() || samples.empty()) return 0;
  const auto e = detail2::encode(samples, preds);
  const auto out = detail2::confusion(e, threshold);
  double total = 0;
  for (int f = 0; f <= LSG_NDOM; ++f)
    for (int ln = 1; ln <= LSG_NDOM; ++ln)
      if (out.coverage[f][ln]) total += (double)out.coverage[f][ln] * ((double)f / (double)ln);
  return e.n ? total / (double)e.n : 0;
}

}  // namespace attribution

// ---- pkg/slo: distribution summaries --------------------------------------------------
namespace slo {
struct Percentiles { double TTFTP50 = 0, TTFTP95 = 0, TTFTP99 = 0, TokensPerSP50 = 0, TokensPerSP95 = 0, RetrievalP95MS = 0; };   // calculator.go:32-40
struct RetrievalBreakdown { double VectorDBMS = 0, NetworkMS = 0, DNSMS = 0; };
struct Snapshot { double TTFTMs = 0, TokensPerS = 0; RetrievalBreakdown Retrieval; };

// calculator.go:121-149
inline double quantile(const std::vector<double> &values, double q) {
  double out = 0;
  Context::check(lsg_segment_quantiles(Context::get(), values.data(), nullptr, values.size(), 1, &q, 1, &out, nullptr));
  return out;
}
inline double nonNegative(double v) { return v < 0 ? 0 : v; }   // calculator.go:151-156
// calculator.go:90-93: negative components count as 0
inline double TotalRetrievalMS(const RetrievalBreakdown &r) { return nonNegative(r.VectorDBMS) + nonNegative(r.NetworkMS) + nonNegative(r.DNSMS); }
// calculator.go:99-118: the three value lists go down as three segments of one call
inline Percentiles Aggregate(const std::vector<Snapshot> &items) {
  if (items.empty()) return Percentiles{};   // calculator.go:96-98
  std::vector<double> v;
  std::vector<uint32_t> seg;
  for (uint32_t k = 0; k < 3; ++k)
    for (auto &it : items) {
      v.push_back(k == 0 ? nonNegative(it.TTFTMs) : k == 1 ? nonNegative(it.TokensPerS) : TotalRetrievalMS(it.Retrieval));
      seg.push_back(k);
    }
  const double q[3] = {0.50, 0.95, 0.99};
  double out[9] = {};
  Context::check(lsg_segment_quantiles(Context::get(), v.data(), seg.data(), v.size(), 3, q, 3, out, nullptr));
  return Percentiles{out[0], out[1], out[2], out[3], out[4], out[7]};
}
}  // namespace slo
}  // namespace llmslo
