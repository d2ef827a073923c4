"""ctypes / numpy mirrors of include/llmslo_gpu.h.  Field order and sizes must match the header;
tests/test_abi.py checks the sizes against the compiled library (lsg_abi_sizes)."""
import ctypes as C

import numpy as np

NSIG = 12
NDOM = 8
SIG_UNSUPPORTED = 15
TRACE_BITS = 28
TS_ZERO = -(1 << 63)
NONE = 0xFFFFFFFF
MAX_FANOUT = 32
MAX_ACTUAL = 16
MAX_WINDOW_NS = (1 << 53) - 1
SKETCH_BINS = 16

TIER_NAMES = ["", "trace_id_exact", "pod_pid_100ms", "pod_conn_250ms", "service_node_500ms"]  # dns.go:59-72
TIER_CONF = [0.0, 1.0, 0.9, 0.8, 0.65]

# DefaultLikelihoods literal order, bayesian.go:69-189
SIGNALS = ["dns_latency_ms", "tcp_retransmits_total", "runqueue_delay_ms", "connect_latency_ms", "tls_handshake_ms",
           "cpu_steal_pct", "cfs_throttled_ms", "mem_reclaim_latency_ms", "disk_io_latency_ms", "syscall_latency_ms",
           "connect_errors_total", "tls_handshake_fail_total"]
SIGNAL_INDEX = {s: i for i, s in enumerate(SIGNALS)}
# AllDomains(), bayesian.go:23-34
DOMAINS = ["network_dns", "network_egress", "cpu_throttle", "memory_pressure", "provider_throttle", "provider_error",
           "retrieval_backend", "unknown"]
DOMAIN_INDEX = {d: i for i, d in enumerate(DOMAINS)}

EVENT_DTYPE = np.dtype([("ts_ns", "<i8"), ("value", "<f8"), ("ident", "<u4"), ("pid", "<u4"), ("conn", "<u4"),
                        ("trace_sig", "<u4")])
SPAN_DTYPE = np.dtype([("ts_ns", "<i8"), ("ident", "<u4"), ("pid", "<u4"), ("conn", "<u4"), ("trace", "<u4"),
                       ("incident", "<u4"), ("reserved", "<u4")])
IDENT_DTYPE = np.dtype([("pod", "<u4"), ("node", "<u4"), ("service", "<u4"), ("grp", "<u4")])
FEATURE_DTYPE = np.dtype([("count", "<u4"), ("sketch", "<u4", (SKETCH_BINS,)), ("_pad", "<u4"), ("sum", "<f8")])
EVENT24_DTYPE = np.dtype([("value", "<f8"), ("a", "<u8"), ("b", "<u8")])   # lsg_event24: the 24-byte upload form
E24_TS_BITS, E24_IDENT_BITS, E24_PID_BITS, E24_TRACE_BITS, E24_CONN_BITS = 42, 14, 22, 26, 20
assert EVENT_DTYPE.itemsize == 32 and SPAN_DTYPE.itemsize == 32 and IDENT_DTYPE.itemsize == 16 and EVENT24_DTYPE.itemsize == 24


class Debug(C.Structure):
    _fields_ = [("unmatched", C.c_uint64), ("low_confidence", C.c_uint64), ("fanout_dropped", C.c_uint64),
                ("unsupported_type", C.c_uint64)]

    def as_tuple(self):
        return (self.unmatched, self.low_confidence, self.fanout_dropped, self.unsupported_type)


class Eval(C.Structure):
    _fields_ = [("tp", C.c_uint64), ("fp", C.c_uint64), ("fn", C.c_uint64), ("tn", C.c_uint64),
                ("tier_correct", C.c_uint64), ("tier_comparable", C.c_uint64),
                ("predicted_by_tier", C.c_uint64 * 5), ("matched_by_tier", C.c_uint64 * 5)]


class Model(C.Structure):
    _fields_ = [("prior", C.c_double * NDOM), ("lik", (C.c_double * NDOM) * NSIG), ("thresh", C.c_double * NSIG),
                ("log_extra", C.c_double * NDOM), ("sig_mask", C.c_uint32), ("reserved", C.c_uint32)]


class ConfusionOut(C.Structure):
    _fields_ = [("matrix", (C.c_uint64 * NDOM) * MAX_ACTUAL), ("n", C.c_uint64), ("correct", C.c_uint64),
                ("partial", C.c_uint64), ("coverage", (C.c_uint64 * (NDOM + 1)) * (NDOM + 1))]


class Counters(C.Structure):
    _fields_ = [("debug", Debug), ("confusion", ConfusionOut), ("events", C.c_uint64), ("spans", C.c_uint64),
                ("incidents", C.c_uint64), ("candidates", C.c_uint64), ("kept", C.c_uint64), ("bad_idents", C.c_uint64),
                ("unsupported_events", C.c_uint64)]


COUNTER_WORDS = C.sizeof(Counters) // 8


class PipelineArgs(C.Structure):
    _fields_ = [("spans", C.c_void_p), ("n_spans", C.c_size_t), ("events", C.c_void_p), ("n_events", C.c_size_t),
                ("idents", C.c_void_p), ("n_idents", C.c_size_t), ("n_incidents", C.c_size_t),
                ("inc_actual", C.c_void_p), ("inc_expected", C.c_void_p), ("inc_expected_n", C.c_void_p),
                ("window_ns", C.c_int64), ("threshold", C.c_double), ("fanout", C.c_int), ("coverage_thr", C.c_double),
                ("cand_event", C.c_void_p), ("cand_tier", C.c_void_p), ("attr_out", C.c_void_p),
                ("conf_out", C.c_void_p), ("inc_signals", C.c_void_p), ("posterior", C.c_void_p),
                ("top1", C.c_void_p), ("counters_dev", C.c_void_p), ("cand_value", C.c_void_p), ("cand_sig", C.c_void_p),
                ("key_hint", C.c_uint32), ("max_grp_hint", C.c_uint32), ("span_ts_lo", C.c_int64), ("span_ts_hi", C.c_int64),
                ("events24", C.c_void_p), ("events24_base", C.c_int64), ("features_out", C.c_void_p)]


class SynthCfg(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_events", C.c_uint64), ("n_spans", C.c_uint64), ("n_pods", C.c_uint32),
                ("pods_per_node", C.c_uint32), ("pods_per_group", C.c_uint32), ("spans_per_incident", C.c_uint32),
                ("multi_fault", C.c_uint32), ("t0_ns", C.c_int64), ("horizon_ns", C.c_int64), ("rank", C.c_uint32),
                ("world", C.c_uint32)]
