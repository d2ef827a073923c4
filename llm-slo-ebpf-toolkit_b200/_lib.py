"""ctypes loader and thin typed wrapper over the C ABI (include/llmslo_gpu.h).

The product path fails loudly: a missing library raises LsgError at first use, a missing CUDA
device makes lsg_create return LSG_ERR_NODEVICE.  Nothing here computes on the CPU.
"""
import ctypes as C
import os
import threading

import numpy as np

from . import _abi as A

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_LOCK = threading.Lock()


class LsgError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libllmslo_gpu: %s (code %d)" % (msg, code))
        self.code = code


def lib_path():
    return os.environ.get("LLMSLO_GPU_LIB", os.path.join(_HERE, "libllmslo_gpu.so"))


def load_library():
    """dlopen libllmslo_gpu.so (built in-tree by __graft_entry__.build()) and declare prototypes."""
    global _LIB
    with _LOCK:
        if _LIB is not None:
            return _LIB
        path = lib_path()
        if not os.path.exists(path):
            raise LsgError(-5, "%s not found: run `python -c 'import __graft_entry__ as g; g.build()'`; "
                               "there is no CPU fallback" % path)
        lib = C.CDLL(path)
        vp, sz, i64, dbl, cint = C.c_void_p, C.c_size_t, C.c_int64, C.c_double, C.c_int
        P = C.POINTER
        lib.lsg_create.argtypes = [cint, vp, P(vp)]
        lib.lsg_destroy.argtypes = [vp]
        lib.lsg_destroy.restype = None
        lib.lsg_last_error.argtypes = [vp]
        lib.lsg_last_error.restype = C.c_char_p
        lib.lsg_version.restype = C.c_char_p
        lib.lsg_launch_count.argtypes = [vp]
        lib.lsg_launch_count.restype = C.c_uint64
        lib.lsg_set_join_path.argtypes = [vp, cint]
        for name in ("lsg_validate_groups", "lsg_validate_groups_dev"):
            getattr(lib, name).argtypes = [vp, vp, sz, vp, sz, vp, sz, P(C.c_uint64)]
        lib.lsg_multi_create.argtypes = [P(cint), cint, P(vp)]
        lib.lsg_multi_destroy.argtypes = [vp]
        lib.lsg_multi_destroy.restype = None
        lib.lsg_multi_last_error.argtypes = [vp]
        lib.lsg_multi_last_error.restype = C.c_char_p
        lib.lsg_multi_size.argtypes = [vp]
        lib.lsg_multi_ctx.argtypes = [vp, cint]
        lib.lsg_multi_ctx.restype = vp
        lib.lsg_multi_pipeline.argtypes = [vp, P(A.PipelineArgs), P(A.Counters), P(A.Counters)]
        lib.lsg_pack_events24.argtypes = [vp, sz, i64, vp]
        lib.lsg_pack_events24.restype = C.c_longlong
        lib.lsg_unpack_events24.argtypes = [vp, sz, i64, vp]
        lib.lsg_widen_events24_dev.argtypes = [vp, vp, sz, i64, vp]
        lib.lsg_workspace_bytes.argtypes = [vp]
        lib.lsg_workspace_bytes.restype = sz
        lib.lsg_default_model.argtypes = [P(A.Model)]
        lib.lsg_default_model.restype = None
        lib.lsg_set_model.argtypes = [vp, P(A.Model)]
        for name in ("lsg_match_pairs", "lsg_match_pairs_dev"):
            getattr(lib, name).argtypes = [vp, vp, vp, sz, vp, sz, i64, dbl, vp, vp, vp, vp, P(A.Eval)]
        for name in ("lsg_process_batch", "lsg_process_batch_dev"):
            getattr(lib, name).argtypes = [vp, vp, sz, vp, sz, vp, sz, i64, dbl, cint, vp, vp, vp, vp, vp, vp, vp,
                                           P(A.Debug)]
        lib.lsg_sort_events_dev.argtypes = [vp, vp, sz, vp, sz, vp, vp, vp]
        for name in ("lsg_incident_features", "lsg_incident_features_dev"):
            getattr(lib, name).argtypes = [vp, vp, vp, sz, sz, vp, vp]
        for name in ("lsg_attribute", "lsg_attribute_dev"):
            getattr(lib, name).argtypes = [vp, vp, vp, vp, sz, vp, vp, vp]
        for name in ("lsg_confusion", "lsg_confusion_dev"):
            getattr(lib, name).argtypes = [vp, vp, vp, vp, vp, vp, dbl, dbl, sz, P(A.ConfusionOut)]
        lib.lsg_pipeline_dev.argtypes = [vp, P(A.PipelineArgs)]
        lib.lsg_synth_n_idents.argtypes = [P(A.SynthCfg)]
        lib.lsg_synth_n_idents.restype = sz
        lib.lsg_synth_n_incidents.argtypes = [P(A.SynthCfg)]
        lib.lsg_synth_n_incidents.restype = sz
        lib.lsg_synth_dev.argtypes = [vp, P(A.SynthCfg), vp, vp, vp, vp, vp, vp]
        lib.lsg_synth_host.argtypes = [P(A.SynthCfg), vp, vp, vp, vp, vp, vp]
        lib.lsg_pipeline.argtypes = [vp, P(A.PipelineArgs), P(A.Counters)]
        lib.lsg_last_sort_info.argtypes = [vp, P(C.c_uint32), P(C.c_uint32), P(C.c_uint32)]
        lib.lsg_pipeline_submit.argtypes = [vp, P(A.PipelineArgs), cint]
        lib.lsg_pipeline_wait.argtypes = [vp, cint, P(A.Counters)]
        lib.lsg_profile_enable.argtypes = [vp, cint]
        lib.lsg_profile_name.argtypes = [cint]
        lib.lsg_profile_name.restype = C.c_char_p
        lib.lsg_profile_read.argtypes = [vp, P(C.c_double), P(C.c_uint64), cint, cint]
        for name in ("lsg_decode_ringbuf", "lsg_decode_ringbuf_dev"):
            getattr(lib, name).argtypes = [vp, vp, sz, C.c_uint32, C.c_uint32, i64, vp, vp, vp]
        for name in ("lsg_signal_status", "lsg_signal_status_dev"):
            getattr(lib, name).argtypes = [vp, vp, sz, vp]
        for name in ("lsg_retry_storm", "lsg_retry_storm_dev"):
            getattr(lib, name).argtypes = [vp, vp, sz, vp, sz, i64, cint, vp, vp]
        for name in ("lsg_segment_quantiles", "lsg_segment_quantiles_dev"):
            getattr(lib, name).argtypes = [vp, vp, vp, sz, sz, vp, sz, vp, vp]
        lib.lsg_incident_signals_static.argtypes = [vp, vp, vp, sz, sz, vp, C.c_char_p, sz]
        lib.lsg_host_alloc.argtypes = [sz, cint, P(vp)]
        lib.lsg_host_free.argtypes = [vp]
        lib.lsg_host_free.restype = None
        lib.lsg_abi_sizes.argtypes = [P(C.c_size_t), cint]
        lib.lsg_abi_sizes.restype = cint
        _LIB = lib
        return lib


def _ptr(a):
    """host numpy array (or None) -> void*"""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"], "array must be C-contiguous"
    return a.ctypes.data


def _as(a, dtype):
    if a is None:
        return None
    return np.ascontiguousarray(a, dtype=dtype)


class Context:
    """One lsg_ctx: one device, one stream.  Not thread-safe (as the C ABI says)."""

    def __init__(self, device=0, stream=None):
        self._lib = load_library()
        h = C.c_void_p()
        rc = self._lib.lsg_create(int(device), C.c_void_p(stream) if stream else None, C.byref(h))
        if rc != 0:
            raise LsgError(rc, "lsg_create failed: no usable CUDA device %d (there is no CPU path)" % device)
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            self._lib.lsg_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise LsgError(rc, (self._lib.lsg_last_error(self._h) or b"?").decode())

    @property
    def handle(self):
        return self._h

    def launch_count(self):
        return int(self._lib.lsg_launch_count(self._h))

    def workspace_bytes(self):
        return int(self._lib.lsg_workspace_bytes(self._h))

    def set_model(self, model):
        self._check(self._lib.lsg_set_model(self._h, C.byref(model)))

    # ---- host-buffer calls ---------------------------------------------------------------
    def match_pairs(self, spans, events, idents, window_ns, threshold, expected_match=None, expected_tier=None):
        spans, events, idents = _as(spans, A.SPAN_DTYPE), _as(events, A.EVENT_DTYPE), _as(idents, A.IDENT_DTYPE)
        n = len(spans)
        assert len(events) == n
        em, et = _as(expected_match, np.uint8), _as(expected_tier, np.uint8)
        tier = np.empty(n, np.uint8)
        pred = np.empty(n, np.uint8)
        ev = A.Eval()
        self._check(self._lib.lsg_match_pairs(self._h, _ptr(spans), _ptr(events), n, _ptr(idents), len(idents),
                                              int(window_ns), float(threshold), _ptr(em), _ptr(et), _ptr(tier),
                                              _ptr(pred), C.byref(ev)))
        return tier, pred, ev

    def process_batch(self, spans, events, idents, window_ns=0, threshold=0.0, fanout=0, attr_base=None):
        spans, events, idents = _as(spans, A.SPAN_DTYPE), _as(events, A.EVENT_DTYPE), _as(idents, A.IDENT_DTYPE)
        ns, ne = len(spans), len(events)
        f = fanout if fanout > 0 else 3
        base = _as(attr_base, np.float64)
        out = {
            "cand_event": np.empty((ns, f), np.uint32), "cand_tier": np.empty((ns, f), np.uint8),
            "attr": np.empty((ns, A.NSIG), np.float64), "conf": np.empty(ns, np.float64),
            "retrieval": np.empty(ns, np.float64), "n_cand": np.empty(ns, np.uint32),
        }
        dbg = A.Debug()
        self._check(self._lib.lsg_process_batch(self._h, _ptr(spans), ns, _ptr(events), ne, _ptr(idents), len(idents),
                                                int(window_ns), float(threshold), int(fanout), _ptr(base),
                                                _ptr(out["cand_event"]), _ptr(out["cand_tier"]), _ptr(out["attr"]),
                                                _ptr(out["conf"]), _ptr(out["retrieval"]), _ptr(out["n_cand"]),
                                                C.byref(dbg)))
        out["debug"] = dbg
        return out

    def incident_features(self, span_attr, span_incident, n_incidents, want_features=False):
        span_attr = _as(span_attr, np.float64)
        span_incident = _as(span_incident, np.uint32)
        ns = len(span_incident)
        sig = np.empty((n_incidents, A.NSIG), np.float64)
        feat = np.zeros((n_incidents, A.NSIG), A.FEATURE_DTYPE) if want_features else None
        self._check(self._lib.lsg_incident_features(self._h, _ptr(span_attr), _ptr(span_incident), ns, n_incidents,
                                                    _ptr(sig), _ptr(feat)))
        return (sig, feat) if want_features else sig

    def attribute(self, signals, has_signals=None, rule_label=None):
        signals = _as(signals, np.float64).reshape(-1, A.NSIG)
        n = len(signals)
        hs, rl = _as(has_signals, np.uint8), _as(rule_label, np.uint8)
        post = np.empty((n, A.NDOM), np.float64)
        top1 = np.empty(n, np.uint8)
        elev = np.empty(n, np.uint16)
        self._check(self._lib.lsg_attribute(self._h, _ptr(signals), _ptr(hs), _ptr(rl), n, _ptr(post), _ptr(top1),
                                            _ptr(elev)))
        return post, top1, elev

    def confusion(self, actual, predicted, expected, expected_n, posterior, coverage_thr, hyp_floor=0.0):
        actual, predicted = _as(actual, np.uint8), _as(predicted, np.uint8)
        expected, expected_n = _as(expected, np.uint32), _as(expected_n, np.uint8)
        posterior = _as(posterior, np.float64)
        out = A.ConfusionOut()
        self._check(self._lib.lsg_confusion(self._h, _ptr(actual), _ptr(predicted), _ptr(expected), _ptr(expected_n),
                                            _ptr(posterior), float(coverage_thr), float(hyp_floor), len(actual),
                                            C.byref(out)))
        return out

    def decode_ringbuf(self, raw, ident=0, trace=0, ts_offset_ns=0):
        """raw: bytes / uint8 array of 40-byte llm_slo_event records -> (events, status, errno)"""
        raw = np.frombuffer(raw, dtype=np.uint8) if isinstance(raw, (bytes, bytearray, memoryview)) else _as(raw, np.uint8)
        if raw.size % 40:
            raise ValueError("decode bpf event: %d bytes is not a multiple of the 40-byte record" % raw.size)
        n = raw.size // 40
        ev = np.zeros(n, A.EVENT_DTYPE)
        st = np.zeros(n, np.uint8)
        er = np.zeros(n, np.int32)
        self._check(self._lib.lsg_decode_ringbuf(self._h, _ptr(np.ascontiguousarray(raw)), n, int(ident), int(trace),
                                                 int(ts_offset_ns), _ptr(ev), _ptr(st), _ptr(er)))
        return ev, st, er

    def signal_status(self, events):
        events = _as(events, A.EVENT_DTYPE)
        st = np.zeros(len(events), np.uint8)
        self._check(self._lib.lsg_signal_status(self._h, _ptr(events), len(events), _ptr(st)))
        return st

    def retry_storm(self, events, idents, window_ns, threshold):
        events, idents = _as(events, A.EVENT_DTYPE), _as(idents, A.IDENT_DTYPE)
        cnt = np.zeros(len(events), np.uint32)
        storm = np.zeros(len(events), np.uint8)
        self._check(self._lib.lsg_retry_storm(self._h, _ptr(events), len(events), _ptr(idents), len(idents), int(window_ns),
                                              int(threshold), _ptr(cnt), _ptr(storm)))
        return cnt, storm

    def segment_quantiles(self, values, segment, n_segments, q):
        """quantile(list s, q[j]) for every list and q (calculator.go:121-149) -> (out[n_segments, n_q], counts)"""
        values, q = _as(values, np.float64), _as(q, np.float64)
        segment = _as(segment, np.uint32)
        if segment is not None and len(segment) != len(values):
            raise ValueError("segment and values differ in length")
        out = np.zeros((int(n_segments), len(q)), np.float64)
        cnt = np.zeros(int(n_segments), np.uint64)
        self._check(self._lib.lsg_segment_quantiles(self._h, _ptr(values), _ptr(segment), len(values), int(n_segments),
                                                    _ptr(q), len(q), _ptr(out), _ptr(cnt)))
        return out, cnt

    def last_sort_info(self):
        b, sh, ps = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self._check(self._lib.lsg_last_sort_info(self._h, C.byref(b), C.byref(sh), C.byref(ps)))
        return {"key_bits": b.value, "time_shift": sh.value, "passes": ps.value}

    def validate_groups(self, spans, events, idents):
        """The join-group contract on a batch: (pod conflicts, (service, node) conflicts, trace-id conflicts, bad idents)."""
        spans, events, idents = _as(spans, A.SPAN_DTYPE), _as(events, A.EVENT_DTYPE), _as(idents, A.IDENT_DTYPE)
        out = (C.c_uint64 * 4)()
        self._check(self._lib.lsg_validate_groups(self._h, _ptr(spans), len(spans), _ptr(events), len(events), _ptr(idents),
                                                  len(idents), out))
        return tuple(int(x) for x in out)

    def set_join_path(self, path):
        """0: permutation sort + gathering join (default), 1: bucket partition + slab join, -1: LSG_JOIN_PATH decides;
        2 / 3: path 0 with the sliding-tile / the fixed-chunk form of its tile kernel forced."""
        self._check(self._lib.lsg_set_join_path(self._h, int(path)))

    def profile_enable(self, on=True):
        self._check(self._lib.lsg_profile_enable(self._h, 1 if on else 0))

    def profile_read(self, reset=True):
        """{kernel name: (total ms, intervals)} since the last reset; synchronises the stream."""
        ms = (C.c_double * 32)()
        cnt = (C.c_uint64 * 32)()
        n = self._lib.lsg_profile_read(self._h, ms, cnt, 32, 1 if reset else 0)
        if n < 0:
            self._check(n)
        return {self._lib.lsg_profile_name(i).decode(): (ms[i], int(cnt[i])) for i in range(n)}

    # ---- device-pointer calls (integers are raw device addresses, e.g. torch.Tensor.data_ptr()) ----
    def call_dev(self, name, *args):
        self._check(getattr(self._lib, name)(self._h, *args))


_DEFAULT = None


class HostBuffer:
    """Page-locked host memory from lsg_host_alloc, viewed as a numpy uint8 array (`.array`); `.ptr` is its address."""

    def __init__(self, nbytes, write_combined=False):
        self._lib = load_library()
        p = C.c_void_p()
        rc = self._lib.lsg_host_alloc(max(int(nbytes), 1), 1 if write_combined else 0, C.byref(p))
        if rc != 0 or not p.value:
            raise LsgError(rc, "lsg_host_alloc(%d bytes) failed" % nbytes)
        self.ptr = p.value
        self.nbytes = int(nbytes)
        self.array = np.ctypeslib.as_array((C.c_uint8 * max(self.nbytes, 1)).from_address(self.ptr))[: self.nbytes]

    def fill_from(self, src):
        """copy the bytes of a contiguous numpy array into the buffer (sequential writes: fine for write-combined memory)"""
        b = np.ascontiguousarray(src).view(np.uint8).reshape(-1)
        assert b.size <= self.nbytes
        self.array[: b.size] = b

    def free(self):
        if self.ptr:
            self.array = None
            self._lib.lsg_host_free(self.ptr)
            self.ptr = 0

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def default_context():
    """Process-wide context on device LLMSLO_GPU_DEVICE (default 0), created on first use."""
    global _DEFAULT
    if _DEFAULT is None:
        _DEFAULT = Context(int(os.environ.get("LLMSLO_GPU_DEVICE", "0")))
    return _DEFAULT
