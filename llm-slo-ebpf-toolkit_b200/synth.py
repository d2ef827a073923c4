"""The synthetic workload of SURVEY.md §8(d) (csrc/synth_common.h): generated on the device for the
benchmark, or by the identical host function for tests.  Input production only — not path arithmetic."""
import ctypes as C

import numpy as np

from . import _abi as A
from ._lib import load_library

T0_NS = 1_771_372_800_000_000_000   # 2026-02-18T00:00:00Z, the reference fixtures' epoch
HOUR_NS = 3_600_000_000_000


def config(n_events, n_spans, n_pods=64, pods_per_node=8, pods_per_group=2, spans_per_incident=10, multi_fault=0,
           seed=42, t0_ns=T0_NS, horizon_ns=HOUR_NS, rank=0, world=1):
    return A.SynthCfg(seed, n_events, n_spans, n_pods, pods_per_node, pods_per_group, spans_per_incident, multi_fault,
                      t0_ns, horizon_ns, rank, world)


# BASELINE.json configs
def c2(**kw):
    return config(10_000_000, 1_000_000, 64, 8, 2, 10, 0, **kw)


def c3(**kw):
    return config(1_000_000_000, 50_000_000, 4096, 16, 2, 10, 1, **kw)


def generate_host(cfg):
    lib = load_library()
    n_id, n_inc = lib.lsg_synth_n_idents(C.byref(cfg)), lib.lsg_synth_n_incidents(C.byref(cfg))
    out = {
        "spans": np.zeros(cfg.n_spans, A.SPAN_DTYPE), "events": np.zeros(cfg.n_events, A.EVENT_DTYPE),
        "idents": np.zeros(n_id, A.IDENT_DTYPE), "inc_actual": np.zeros(n_inc, np.uint8),
        "inc_expected": np.zeros(n_inc, np.uint32), "inc_expected_n": np.zeros(n_inc, np.uint8),
    }
    rc = lib.lsg_synth_host(C.byref(cfg), out["spans"].ctypes.data, out["events"].ctypes.data, out["idents"].ctypes.data,
                            out["inc_actual"].ctypes.data, out["inc_expected"].ctypes.data,
                            out["inc_expected_n"].ctypes.data)
    if rc != 0:
        raise ValueError("invalid synthetic workload configuration")
    return out
