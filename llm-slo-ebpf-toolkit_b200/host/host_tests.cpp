// host_tests.cpp — the reference's unit tests for the hot path, transcribed onto the C++ host layer (llmslo.hpp).
// Sources: pkg/correlation/dns_test.go, evaluator_test.go; pkg/otel/processor/ebpfcorrelator/correlator_test.go;
// pkg/attribution/bayesian_test.go, pipeline_test.go, mapper_test.go.  Needs a GPU (there is no CPU path).
#include <chrono>
#include <cstdio>

#include "llmslo.hpp"

using namespace llmslo;
namespace co = llmslo::correlation;
namespace ec = llmslo::ebpfcorrelator;
namespace at = llmslo::attribution;

static int g_fail = 0, g_checks = 0;
#define CHECK(cond)                                                         \
  do {                                                                      \
    ++g_checks;                                                             \
    if (!(cond)) { ++g_fail; printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond); } \
  } while (0)

static const Time NOW(1771372800123456789ll);

static co::SignalRef sig(const char *name, int64_t dt_ms, double value, int64_t pid = 101, const char *trace = "trace-9") {
  co::SignalRef s;
  s.Signal = name; s.TraceID = trace; s.Service = "rag"; s.Node = "node-a"; s.Pod = "pod-a"; s.PID = pid;
  s.Timestamp = NOW.Add(dt_ms * Millisecond); s.Value = value;
  return s;
}

static void TestMatchTiers() {   // dns_test.go:8-87
  co::SpanRef span{"trace-1", "chat", "node-a", "pod-a", 42, "10.0.0.1:1234->10.0.0.2:53/udp", NOW};
  co::SignalRef a; a.Signal = "dns_latency_ms"; a.TraceID = "trace-1"; a.Timestamp = NOW.Add(1 * Second);
  co::SignalRef b; b.Signal = "dns_latency_ms"; b.Pod = "pod-a"; b.PID = 42; b.Timestamp = NOW.Add(50 * Millisecond);
  co::SignalRef c; c.Signal = "dns_latency_ms"; c.Pod = "pod-a"; c.ConnTuple = "10.0.0.1:1234->10.0.0.2:53/udp"; c.Timestamp = NOW.Add(200 * Millisecond);
  co::SignalRef d; d.Signal = "dns_latency_ms"; d.Service = "chat"; d.Node = "node-a"; d.Timestamp = NOW.Add(400 * Millisecond);
  struct { co::SignalRef s; double conf; const char *tier; } tests[] = {
      {a, 1.0, "trace_id_exact"}, {b, 0.9, "pod_pid_100ms"}, {c, 0.8, "pod_conn_250ms"}, {d, 0.65, "service_node_500ms"}};
  for (auto &t : tests) {
    co::Decision dec = co::Match(span, t.s, co::DefaultWindow);
    CHECK(dec.Matched);
    CHECK(dec.Confidence == t.conf);
    CHECK(dec.Tier == t.tier);
  }
  co::SpanRef zero{"t", "", "", "", 0, "", Time()};   // zero Time never matches; Unix 0 does
  co::SignalRef z; z.TraceID = "t"; z.Timestamp = NOW;
  CHECK(!co::Match(zero, z, 0).Matched);
  co::SpanRef unix0{"t", "", "", "", 0, "", Time(0)};
  z.Timestamp = Time(0);
  CHECK(co::Match(unix0, z, 0).Matched);
}

static void TestEvaluateLabeledPairs() {   // evaluator_test.go:10-86
  co::LabeledPair p1, p2, p3;
  p1.CaseID = "tp-trace"; p1.ExpectedMatch = true; p1.ExpectedTier = "trace_id_exact";
  p1.Span.TraceID = "trace-1"; p1.Span.Timestamp = NOW; p1.Signal.Signal = "dns_latency_ms"; p1.Signal.TraceID = "trace-1";
  p1.Signal.Timestamp = NOW.Add(10 * Millisecond);
  p2.CaseID = "fn-low-confidence"; p2.ExpectedMatch = true; p2.ExpectedTier = "service_node_500ms";
  p2.Span.Service = "svc-a"; p2.Span.Node = "node-a"; p2.Span.Timestamp = NOW; p2.Signal.Signal = "dns_latency_ms";
  p2.Signal.Service = "svc-a"; p2.Signal.Node = "node-a"; p2.Signal.Timestamp = NOW.Add(100 * Millisecond);
  p3.CaseID = "tn-unmatched"; p3.ExpectedMatch = false;
  p3.Span.Service = "svc-a"; p3.Span.Node = "node-a"; p3.Span.Timestamp = NOW; p3.Signal.Signal = "dns_latency_ms";
  p3.Signal.Service = "svc-b"; p3.Signal.Node = "node-b"; p3.Signal.Timestamp = NOW.Add(10 * Millisecond);
  auto r = co::EvaluateLabeledPairs({p1, p2, p3}, 2 * Second, 0.7);
  CHECK(r.second.size() == 3);
  CHECK(r.first.TruePositive == 1 && r.first.FalseNegative == 1 && r.first.TrueNegative == 1);
  CHECK(r.first.Precision == 1.0);
  CHECK(r.first.Recall == 0.5);
  co::EvalReport rep;
  rep.Precision = 0.91; rep.Recall = 0.86;
  CHECK(co::EvaluateGate(rep, 0.9, 0.85).Pass);
  CHECK(!co::EvaluateGate(rep, 0.95, 0.85).Pass);
  CHECK(co::EvaluateLabeledPairs({}, 0, 0).second.empty());
}

static void TestCorrelator() {   // correlator_test.go:38-161, 206-324
  co::SpanRef span{"trace-9", "rag", "node-a", "pod-a", 101, "10.0.0.2:42424->10.0.0.53:443/tcp", NOW};
  auto res = ec::New().EnrichAttributes({}, span, {sig("dns_latency_ms", 10, 120), sig("tcp_retransmits_total", 15, 5),
                                                   sig("runqueue_delay_ms", 20, 22), sig("connect_latency_ms", 22, 180)});
  CHECK(res.Candidates.size() == 3);
  CHECK(res.Debug.FanoutDropped == 1);
  CHECK(res.Attributes.count("llm.ebpf.dns.latency_ms") && res.Attributes.count("llm.ebpf.tcp.retransmits"));
  CHECK(res.Attributes.count("llm.ebpf.sched.runqueue_delay_ms"));
  CHECK(!res.Attributes.count("llm.ebpf.net.connect_latency_ms"));   // the farthest in time is dropped

  co::SpanRef s2; s2.Service = "svc-a"; s2.Node = "node-a"; s2.Pod = "pod-a"; s2.Timestamp = NOW;
  co::SignalRef g1; g1.Signal = "dns_latency_ms"; g1.Service = "svc-a"; g1.Node = "node-a"; g1.Pod = "pod-a"; g1.Timestamp = NOW.Add(300 * Millisecond); g1.Value = 99;
  co::SignalRef g2; g2.Signal = "cpu_steal_pct"; g2.Service = "svc-b"; g2.Node = "node-z"; g2.Pod = "pod-z"; g2.Timestamp = NOW.Add(20 * Millisecond); g2.Value = 9;
  co::SignalRef g3; g3.Signal = "unknown_signal"; g3.TraceID = "trace-a"; g3.Timestamp = NOW; g3.Value = 1;
  res = ec::New().EnrichAttributes({}, s2, {g1, g2, g3});
  CHECK(res.Candidates.empty());
  CHECK(res.Debug.LowConfidence == 1 && res.Debug.Unmatched == 1 && res.Debug.UnsupportedType == 1);

  ec::SpanRecord r1, r2;
  r1.TraceID = "trace-1"; r1.Service = "rag"; r1.Node = "node-a"; r1.Pod = "pod-a"; r1.PID = 1001; r1.Timestamp = NOW;
  r2 = r1; r2.TraceID = "trace-2"; r2.PID = 1002;
  co::SignalRef d1 = sig("dns_latency_ms", 10, 88, 1001, "trace-1");
  co::SignalRef d2; d2.Signal = "unknown_signal"; d2.TraceID = "trace-1"; d2.Timestamp = NOW; d2.Value = 1;
  auto out = ec::New().ProcessBatch({r1, r2}, {d1, d2});
  CHECK(out.Spans.size() == 2);
  CHECK(out.Spans[0].Attributes["llm.ebpf.dns.latency_ms"] == 88);
  CHECK(out.Spans[0].Attributes["llm.ebpf.retrieval.kernel_attributed_ms"] == 88);
  CHECK(!out.Spans[1].Attributes.count("llm.ebpf.dns.latency_ms"));
  CHECK(out.Debug.UnsupportedType == 2);
  CHECK(ec::New().ProcessBatch({}, {d1}).Spans.empty());

  ec::SpanRecord r9 = r1;
  r9.TraceID = "trace-9"; r9.PID = 222;
  out = ec::New().ProcessBatch({r9}, {sig("dns_latency_ms", 5, 10, 222), sig("connect_latency_ms", 6, 11, 222),
                                      sig("tls_handshake_ms", 7, 12, 222), sig("runqueue_delay_ms", 8, 13, 222)});
  CHECK(out.Debug.FanoutDropped == 1);
  CHECK(out.Spans[0].Attributes["llm.ebpf.retrieval.kernel_attributed_ms"] == 10 + 11 + 12);
  ec::AttrMap m = {{"llm.ebpf.dns.latency_ms", 12.5}, {"llm.ebpf.net.connect_latency_ms", 18.0}, {"llm.ebpf.tls.handshake_ms", 22.0},
                   {"llm.ebpf.cpu.steal_pct", 0.6}};
  CHECK(ec::DecomposeRetrieval(m) == 12.5 + 18.0 + 22.0);
  ec::AttrMap m2 = {{"llm.ebpf.cpu.steal_pct", 2.0}};
  CHECK(ec::DecomposeRetrieval(m2) == 0 && !m2.count(ec::AttrRetrievalKernelMS));
}

static void TestBayesian() {   // bayesian_test.go:9-195
  at::BayesianAttributor ba = at::NewBayesianAttributor();
  double sum = 0;
  for (auto &p : ba.Attribute({{"dns_latency_ms", 220}, {"connect_latency_ms", 130}})) sum += p.Posterior;
  CHECK(std::fabs(sum - 1.0) <= 1e-6);
  struct { at::SignalMap s; const char *want; } cases[] = {
      {{{"dns_latency_ms", 220}, {"connect_latency_ms", 130}, {"tcp_retransmits_total", 0.2}, {"runqueue_delay_ms", 4}, {"cpu_steal_pct", 0.6}}, "network_dns"},
      {{{"runqueue_delay_ms", 28}, {"cpu_steal_pct", 9}, {"cfs_throttled_ms", 170}, {"dns_latency_ms", 12}}, "cpu_throttle"},
      {{{"cfs_throttled_ms", 90}, {"mem_reclaim_latency_ms", 25}, {"disk_io_latency_ms", 60}, {"runqueue_delay_ms", 14}, {"dns_latency_ms", 12}}, "memory_pressure"},
      {{{"connect_latency_ms", 95}, {"tls_handshake_ms", 90}, {"syscall_latency_ms", 300}, {"connect_errors_total", 2}, {"dns_latency_ms", 12}}, "provider_throttle"}};
  for (auto &c : cases) CHECK(ba.Attribute(c.s)[0].Domain == c.want);
  auto p0 = ba.Attribute(cases[0].s);
  CHECK(p0[0].Posterior >= 0.3);
  CHECK(std::fabs(p0[0].Posterior - 0.9332699048386807) <= 1e-6 * 0.9332699048386807);   // mf-01, SURVEY.md §4
  CHECK((p0[0].Evidence == std::vector<std::string>{"connect_latency_ms", "dns_latency_ms"}));
  CHECK(at::AllDomains().size() == 8);
  for (auto &kv : ba.Priors) CHECK(kv.second == 1.0 / 8.0);
  at::FaultSample s;
  s.IncidentID = "i"; s.FaultLabel = "dns_latency"; s.Confidence = 0.9;   // no Signals -> rule path, bayesian.go:318-320
  auto a = ba.AttributeSample(s);
  CHECK(a.PredictedFaultDomain == "network_dns" && a.Confidence == 0.9 && a.FaultHypotheses.empty());
  CHECK(at::MapFaultLabel("network_partition") == "network_egress" && at::MapFaultLabel("whatever") == "unknown");
}

static at::IncidentAttribution pred(const char *dom, std::vector<std::pair<const char *, double>> hyps = {}) {
  at::IncidentAttribution a;
  a.PredictedFaultDomain = dom;
  for (auto &h : hyps) a.FaultHypotheses.push_back(at::FaultHypothesis{h.first, h.second, {}});
  return a;
}
static at::FaultSample sample(const char *label, const char *expected = "", std::vector<std::string> exps = {}) {
  at::FaultSample s;
  s.FaultLabel = label; s.ExpectedDomain = expected; s.ExpectedDomains = std::move(exps);
  return s;
}

static void TestPipeline() {   // pipeline_test.go:46-175
  auto m = at::BuildConfusionMatrix({sample("provider_throttle", "network_dns")}, {pred("provider_throttle")});
  CHECK(m.size() == 1 && (m[at::MatrixKey{"network_dns", "provider_throttle"}] == 1));
  CHECK(at::Accuracy({sample("provider_throttle"), sample("dns_latency")}, {pred("provider_throttle"), pred("provider_throttle")}) == 0.5);
  CHECK(at::PartialAccuracy({sample("dns_latency", "", {"network_dns"}), sample("cpu_throttle", "", {"cpu_throttle"})},
                            {pred("network_dns"), pred("cpu_throttle")}) == 1.0);
  CHECK(at::PartialAccuracy({sample("", "", {"network_dns", "cpu_throttle"}), sample("", "", {"provider_throttle", "memory_pressure"})},
                            {pred("cpu_throttle"), pred("network_dns")}) == 0.5);
  std::vector<at::FaultSample> s = {sample("", "", {"network_dns", "cpu_throttle"})};
  CHECK(at::CoverageAccuracy(s, {pred("network_dns", {{"network_dns", 0.5}, {"cpu_throttle", 0.3}})}, 0.1) == 1.0);
  CHECK(at::CoverageAccuracy(s, {pred("network_dns", {{"network_dns", 0.5}, {"memory_pressure", 0.3}})}, 0.1) == 0.5);
  CHECK(at::CoverageAccuracy(s, {pred("network_dns", {{"network_dns", 0.5}, {"cpu_throttle", 0.02}})}, 0.05) == 0.5);
  CHECK(at::Accuracy({}, {}) == 0);
  CHECK(at::BuildAttributions({sample("dns_latency")}, " RULE ")[0].PredictedFaultDomain == "network_dns");
}

// pkg/slo/calculator_test.go:34-52 (TestAggregate) + quantile edge cases of calculator.go:121-149
static void TestAggregate() {
  namespace sl = llmslo::slo;
  std::vector<sl::Snapshot> items = {{100, 50, {20, 10, 5}}, {200, 30, {40, 15, 10}}, {300, 10, {60, 25, 15}}};
  const auto out = sl::Aggregate(items);
  CHECK(out.TTFTP50 == 200);
  CHECK(out.TokensPerSP50 == 30);
  CHECK(out.RetrievalP95MS > 0);
  const double frac = 0.95 * 2.0 - 1.0;
  CHECK(out.RetrievalP95MS == 65.0 * (1 - frac) + 100.0 * frac);
  CHECK(sl::quantile({}, 0.5) == 0);
  CHECK(sl::quantile({7.0}, 0.99) == 7.0);
  CHECK(sl::quantile({3.0, 1.0, 2.0}, -1) == 1.0);
  CHECK(sl::quantile({3.0, 1.0, 2.0}, 5) == 3.0);
  CHECK(sl::Aggregate({}).TTFTP95 == 0);
}

// --time N: what a caller of the unchanged string signatures pays end to end (ProcessBatch on N spans and 10 N signals
// held as strings): interning, the library call with pageable host buffers, rebuilding the attribute maps.
static int time_strings(size_t n_spans) {
  using namespace llmslo;
  std::vector<ebpfcorrelator::SpanRecord> spans(n_spans);
  std::vector<correlation::SignalRef> signals;
  signals.reserve(n_spans * 10);
  const char *names[9] = {"dns_latency_ms", "tcp_retransmits_total", "runqueue_delay_ms", "connect_latency_ms", "tls_handshake_ms",
                          "cpu_steal_pct", "mem_reclaim_latency_ms", "disk_io_latency_ms", "syscall_latency_ms"};
  uint64_t x = 42;
  auto rnd = [&]() { x ^= x << 13; x ^= x >> 7; x ^= x << 17; return x; };
  const int64_t t0 = 1771372800000000000ll, horizon = (int64_t)n_spans * 3600000ll;   // C2's density: 1 M spans per hour
  for (size_t i = 0; i < n_spans; ++i) {
    auto &s = spans[i];
    const unsigned pod = (unsigned)(rnd() % 64);
    s.TraceID = "trace-" + std::to_string(i);
    s.Pod = "pod-" + std::to_string(pod); s.Node = "node-" + std::to_string(pod / 8); s.Service = "svc-" + std::to_string(pod / 2 % 4);
    s.PID = 1000 + (int)(rnd() % 4); s.ConnTuple = "10.0.0." + std::to_string(rnd() % 8) + ":443";
    s.Timestamp = Time(t0 + (int64_t)(rnd() % (uint64_t)horizon));
    for (int k = 0; k < 10; ++k) {
      correlation::SignalRef g;
      g.Signal = names[rnd() % 9];
      const unsigned kind = (unsigned)(rnd() % 100);
      g.TraceID = kind < 50 ? s.TraceID : "";
      g.Pod = s.Pod; g.Node = s.Node; g.Service = s.Service; g.PID = kind < 80 ? s.PID : s.PID + 100;
      g.ConnTuple = kind >= 50 && kind < 80 ? "" : s.ConnTuple;
      g.Timestamp = s.Timestamp.Add((int64_t)(rnd() % 200000000ull) - 100000000ll);
      g.Value = 1.0 + (double)(rnd() % 1000) / 4.0;
      signals.push_back(std::move(g));
    }
  }
  auto c = ebpfcorrelator::New();
  double best = 1e300;
  ebpfcorrelator::Correlator::Timing bt{0, 0, 0};
  size_t kept = 0;
  for (int rep = 0; rep < 3; ++rep) {
    const auto a = std::chrono::steady_clock::now();
    auto out = c.ProcessBatch(spans, signals);
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - a).count();
    if (ms < best) { best = ms; bt = ebpfcorrelator::Correlator::last_timing(); }
    kept = 0;
    for (auto &s : out.Spans) kept += s.Candidates.size();
  }
  printf("{\"spans\": %zu, \"signals\": %zu, \"kept\": %zu, \"total_ms\": %.2f, \"intern_ms\": %.2f, \"library_ms\": %.2f, "
         "\"rebuild_ms\": %.2f, \"events_per_s\": %.0f, \"note\": \"ProcessBatch from std::string records through llmslo.hpp, one host "
         "thread, pageable buffers, best of 3\"}\n",
         spans.size(), signals.size(), kept, best, bt.intern_ms, bt.library_ms, bt.rebuild_ms, signals.size() / (best / 1e3));
  return 0;
}

int main(int argc, char **argv) {
  if (argc >= 3 && std::string(argv[1]) == "--time") {
    try {
      return time_strings((size_t)atoll(argv[2]));
    } catch (const std::exception &e) {
      printf("EXCEPTION %s\n", e.what());
      return 2;
    }
  }
  try {
    TestMatchTiers();
    TestEvaluateLabeledPairs();
    TestCorrelator();
    TestBayesian();
    TestPipeline();
    TestAggregate();
  } catch (const std::exception &e) {
    printf("EXCEPTION %s\n", e.what());
    return 2;
  }
  printf("%s: %d checks, %d failed\n", g_fail ? "FAILED" : "ok", g_checks, g_fail);
  return g_fail ? 1 : 0;
}
