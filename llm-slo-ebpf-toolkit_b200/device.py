"""Device-resident workloads: torch tensors as HBM buffers for the `_dev` entry points of the C ABI.
torch is plumbing here (allocation, streams, torch.distributed); every kernel that runs is ours."""
import ctypes as C

import numpy as np
import torch

from . import _abi as A
from ._lib import Context, load_library


def _u8(n, device):
    return torch.empty(max(int(n), 1), dtype=torch.uint8, device=device)


class DeviceWorkload:
    """Inputs and outputs of lsg_pipeline_dev in HBM for one synthetic configuration."""

    def __init__(self, ctx: Context, cfg: A.SynthCfg, fanout=3, window_ns=0, threshold=0.0, coverage_thr=0.10,
                 generate=True):
        self.ctx, self.cfg, self.fanout = ctx, cfg, fanout
        lib = load_library()
        dev = torch.device("cuda", ctx.device)
        self.n_events, self.n_spans = int(cfg.n_events), int(cfg.n_spans)
        self.n_idents = int(lib.lsg_synth_n_idents(C.byref(cfg)))
        self.n_incidents = int(lib.lsg_synth_n_incidents(C.byref(cfg)))
        self.events = _u8(self.n_events * 32, dev)
        self.spans = _u8(self.n_spans * 32, dev)
        self.idents = _u8(self.n_idents * 16, dev)
        self.inc_actual = _u8(self.n_incidents, dev)
        self.inc_expected = torch.empty(max(self.n_incidents, 1), dtype=torch.int32, device=dev)
        self.inc_expected_n = _u8(self.n_incidents, dev)
        f = fanout if fanout > 0 else 3
        self.cand_event = torch.empty(max(self.n_spans * f, 1), dtype=torch.int32, device=dev)
        self.cand_tier = _u8(self.n_spans * f, dev)
        self.attr = torch.empty(max(self.n_spans * A.NSIG, 1), dtype=torch.float64, device=dev)
        self.conf = torch.empty(max(self.n_spans, 1), dtype=torch.float64, device=dev)
        self.inc_signals = torch.empty(max(self.n_incidents * A.NSIG, 1), dtype=torch.float64, device=dev)
        self.posterior = torch.empty(max(self.n_incidents * A.NDOM, 1), dtype=torch.float64, device=dev)
        self.top1 = _u8(self.n_incidents, dev)
        self.counters = torch.zeros(A.COUNTER_WORDS, dtype=torch.int64, device=dev)
        self.args = A.PipelineArgs(
            self.spans.data_ptr(), self.n_spans, self.events.data_ptr(), self.n_events, self.idents.data_ptr(),
            self.n_idents, self.n_incidents, self.inc_actual.data_ptr(), self.inc_expected.data_ptr(),
            self.inc_expected_n.data_ptr(), int(window_ns), float(threshold), int(fanout), float(coverage_thr),
            self.cand_event.data_ptr(), self.cand_tier.data_ptr(), self.attr.data_ptr(), self.conf.data_ptr(),
            self.inc_signals.data_ptr(), self.posterior.data_ptr(), self.top1.data_ptr(), self.counters.data_ptr())
        if generate:
            self.generate()

    def generate(self):
        self.ctx.call_dev("lsg_synth_dev", C.byref(self.cfg), self.spans.data_ptr(), self.events.data_ptr(),
                          self.idents.data_ptr(), self.inc_actual.data_ptr(), self.inc_expected.data_ptr(),
                          self.inc_expected_n.data_ptr())

    def upload(self, w):
        """Load a host-generated workload (dict of numpy arrays from synth.generate_host)."""
        for name in ("events", "spans", "idents", "inc_actual", "inc_expected", "inc_expected_n"):
            src = torch.from_numpy(np.ascontiguousarray(w[name]).view(np.uint8).copy())
            getattr(self, name).view(torch.uint8)[: src.numel()].copy_(src)
        torch.cuda.synchronize()

    def state_key_range(self, span_ts_lo=None, span_ts_hi=None, max_grp=None):
        """Tell the library the spans' time range and the largest join group (lsg_pipeline_args.key_hint): the step then
        runs without the readback of the key layout.  Defaults: what the generator's configuration guarantees (span
        times in [t0, t0 + horizon), one group per `pods_per_group` pods)."""
        cfg = self.cfg
        self.args.key_hint = 1
        self.args.span_ts_lo = int(cfg.t0_ns if span_ts_lo is None else span_ts_lo)
        self.args.span_ts_hi = int(cfg.t0_ns + cfg.horizon_ns - 1 if span_ts_hi is None else span_ts_hi)
        ppg = max(int(cfg.pods_per_group), 1)
        # the generator numbers groups from 1: identity row // pods_per_group + 1
        self.args.max_grp_hint = int((max(self.n_idents, 1) - 1) // ppg + 1 if max_grp is None else max_grp)

    def step(self):
        """One pass of the hot path over the batch (asynchronous on the context's stream)."""
        self.ctx.call_dev("lsg_pipeline_dev", C.byref(self.args))

    def read_counters(self):
        c = A.Counters()
        host = self.counters.cpu().numpy()
        C.memmove(C.byref(c), host.ctypes.data, C.sizeof(c))
        return c

    def host(self, name, dtype, shape=None):
        arr = getattr(self, name).cpu().numpy().view(dtype)
        return arr.reshape(shape) if shape else arr

    def algorithmic_bytes(self):
        """96 + 152 r + 163 q bytes per event."""
        e, s, i = self.n_events, self.n_spans, self.n_incidents
        return 96 * e + 152 * s + 163 * i
