"""ctypes wrapper over oracle/liboracle.so (the packed-record C++ oracle).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess

import numpy as np

from llm_slo_ebpf_toolkit_b200 import _abi as A

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        vp, sz, i64, dbl, cint = C.c_void_p, C.c_size_t, C.c_int64, C.c_double, C.c_int
        L.orc_match.argtypes = [vp, vp, vp, i64]
        L.orc_match_pairs.argtypes = [vp, vp, sz, vp, i64, dbl, vp, vp, vp, vp, C.POINTER(A.Eval)]
        L.orc_match_pairs.restype = None
        for name in ("orc_process_batch_brute", "orc_process_batch_indexed"):
            f = getattr(L, name)
            f.argtypes = [vp, sz, vp, sz, vp, i64, dbl, cint, vp, vp, vp, vp, vp, vp, vp, C.POINTER(A.Debug), cint]
            f.restype = None
        L.orc_incident_features.argtypes = [vp, vp, sz, sz, vp, vp]
        L.orc_incident_features.restype = None
        L.orc_default_model.argtypes = [C.POINTER(A.Model)]
        L.orc_default_model.restype = None
        L.orc_attribute.argtypes = [C.POINTER(A.Model), vp, vp, vp, sz, vp, vp, vp]
        L.orc_attribute.restype = None
        L.orc_confusion.argtypes = [vp, vp, vp, vp, vp, dbl, dbl, sz, C.POINTER(A.ConfusionOut)]
        L.orc_confusion.restype = None
        L.orc_retry_storm.argtypes = [vp, vp, sz, C.c_uint32, i64, cint, vp, vp]
        L.orc_segment_quantiles.argtypes = [vp, vp, sz, sz, vp, sz, vp, vp]
        L.orc_segment_quantiles.restype = None
        L.orc_retry_storm.restype = None
        L.orc_hw_threads.restype = cint
        L.orc_synth_n_incidents.argtypes = [C.POINTER(A.SynthCfg)]
        L.orc_synth_n_incidents.restype = sz
        L.orc_synth_range.argtypes = [C.POINTER(A.SynthCfg), C.c_uint64, C.c_uint64, vp, C.c_uint64, C.c_uint64, vp, vp, vp, vp,
                                      vp, cint]
        _LIB = L
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data


def _c(a, dt):
    return None if a is None else np.ascontiguousarray(a, dtype=dt)


def match_pairs(spans, events, idents, window_ns, threshold, expected_match=None, expected_tier=None):
    spans, events, idents = _c(spans, A.SPAN_DTYPE), _c(events, A.EVENT_DTYPE), _c(idents, A.IDENT_DTYPE)
    n = len(spans)
    em, et = _c(expected_match, np.uint8), _c(expected_tier, np.uint8)
    tier, pred, ev = np.empty(n, np.uint8), np.empty(n, np.uint8), A.Eval()
    lib().orc_match_pairs(_p(spans), _p(events), n, _p(idents), int(window_ns), float(threshold), _p(em), _p(et),
                          _p(tier), _p(pred), C.byref(ev))
    return tier, pred, ev


def process_batch(spans, events, idents, window_ns=0, threshold=0.0, fanout=0, attr_base=None, indexed=False,
                  nthreads=1):
    spans, events, idents = _c(spans, A.SPAN_DTYPE), _c(events, A.EVENT_DTYPE), _c(idents, A.IDENT_DTYPE)
    ns, ne = len(spans), len(events)
    f = fanout if fanout > 0 else 3
    base = _c(attr_base, np.float64)
    out = {"cand_event": np.empty((ns, f), np.uint32), "cand_tier": np.empty((ns, f), np.uint8),
           "attr": np.empty((ns, A.NSIG), np.float64), "conf": np.empty(ns, np.float64),
           "retrieval": np.empty(ns, np.float64), "n_cand": np.empty(ns, np.uint32)}
    dbg = A.Debug()
    fn = lib().orc_process_batch_indexed if indexed else lib().orc_process_batch_brute
    fn(_p(spans), ns, _p(events), ne, _p(idents), int(window_ns), float(threshold), int(fanout), _p(base),
       _p(out["cand_event"]), _p(out["cand_tier"]), _p(out["attr"]), _p(out["conf"]), _p(out["retrieval"]),
       _p(out["n_cand"]), C.byref(dbg), int(nthreads))
    out["debug"] = dbg
    return out


def incident_features(span_attr, span_incident, n_incidents, want_features=False):
    span_attr, span_incident = _c(span_attr, np.float64), _c(span_incident, np.uint32)
    sig = np.empty((n_incidents, A.NSIG), np.float64)
    feat = np.zeros((n_incidents, A.NSIG), A.FEATURE_DTYPE) if want_features else None
    lib().orc_incident_features(_p(span_attr), _p(span_incident), len(span_incident), n_incidents, _p(sig), _p(feat))
    return (sig, feat) if want_features else sig


def default_model():
    m = A.Model()
    lib().orc_default_model(C.byref(m))
    return m


def attribute(model, signals, has_signals=None, rule_label=None):
    signals = _c(signals, np.float64).reshape(-1, A.NSIG)
    n = len(signals)
    hs, rl = _c(has_signals, np.uint8), _c(rule_label, np.uint8)
    post, top1, elev = np.empty((n, A.NDOM)), np.empty(n, np.uint8), np.empty(n, np.uint16)
    lib().orc_attribute(C.byref(model), _p(signals), _p(hs), _p(rl), n, _p(post), _p(top1), _p(elev))
    return post, top1, elev


def confusion(actual, predicted, expected, expected_n, posterior, coverage_thr, hyp_floor=0.0):
    actual, predicted = _c(actual, np.uint8), _c(predicted, np.uint8)
    expected, expected_n = _c(expected, np.uint32), _c(expected_n, np.uint8)
    posterior = _c(posterior, np.float64)
    out = A.ConfusionOut()
    lib().orc_confusion(_p(actual), _p(predicted), _p(expected), _p(expected_n), _p(posterior), float(coverage_thr),
                        float(hyp_floor), len(actual), C.byref(out))
    return out


def retry_storm(pod, ts, n_pods, window_ns, threshold):
    pod, ts = _c(pod, np.uint32), _c(ts, np.int64)
    cnt, storm = np.empty(len(pod), np.uint32), np.empty(len(pod), np.uint8)
    lib().orc_retry_storm(_p(pod), _p(ts), len(pod), int(n_pods), int(window_ns), int(threshold), _p(cnt), _p(storm))
    return cnt, storm


def segment_quantiles(values, segment, n_segments, q):
    values, q = _c(values, np.float64), _c(q, np.float64)
    segment = None if segment is None else _c(segment, np.uint32)
    out, cnt = np.zeros((n_segments, len(q)), np.float64), np.zeros(n_segments, np.uint64)
    lib().orc_segment_quantiles(_p(values), None if segment is None else _p(segment), len(values), int(n_segments), _p(q),
                                len(q), _p(out), _p(cnt))
    return out, cnt


def hw_threads():
    return lib().orc_hw_threads()


def generate_host(cfg, nthreads=0):
    """The synthetic workload of csrc/synth_common.h generated by the host-only copy inside liboracle.so (same header,
    same bytes as lsg_synth_host / lsg_synth_dev), so that a caller need not map the product library."""
    L = lib()
    n_inc = L.orc_synth_n_incidents(C.byref(cfg))
    out = {"spans": np.zeros(cfg.n_spans, A.SPAN_DTYPE), "events": np.zeros(cfg.n_events, A.EVENT_DTYPE),
           "idents": np.zeros(cfg.n_pods, A.IDENT_DTYPE), "inc_actual": np.zeros(n_inc, np.uint8),
           "inc_expected": np.zeros(n_inc, np.uint32), "inc_expected_n": np.zeros(n_inc, np.uint8)}
    rc = L.orc_synth_range(C.byref(cfg), 0, cfg.n_spans, _p(out["spans"]), 0, cfg.n_events, _p(out["events"]), _p(out["idents"]),
                           _p(out["inc_actual"]), _p(out["inc_expected"]), _p(out["inc_expected_n"]),
                           int(nthreads) if nthreads > 0 else max(1, hw_threads()))
    if rc != 0:
        raise ValueError("invalid synthetic workload configuration")
    return out
