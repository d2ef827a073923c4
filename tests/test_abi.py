"""CPU-side checks of the drop-in boundary: the library builds, loads, exports every symbol the header
declares, its struct layouts match the Python mirrors, and it refuses to work without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import llm_slo_ebpf_toolkit_b200 as pkg
from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(pkg.lib_path()):
        import __graft_entry__ as g
        g.build()
    return pkg.load_library()


def test_header_symbols_exported(lib):
    header = open(os.path.join(ROOT, "include", "llmslo_gpu.h")).read()
    names = sorted(set(re.findall(r"\b(lsg_[a-z0-9_]+)\s*\(", header)))
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), "library does not export %s" % n


def test_struct_sizes_match_mirrors(lib):
    out = (C.c_size_t * 16)()
    m = lib.lsg_abi_sizes(out, 16)
    assert m == 12
    want = [A.EVENT_DTYPE.itemsize, A.SPAN_DTYPE.itemsize, A.IDENT_DTYPE.itemsize, C.sizeof(A.Debug), C.sizeof(A.Eval),
            C.sizeof(A.Model), C.sizeof(A.ConfusionOut), A.FEATURE_DTYPE.itemsize, C.sizeof(A.Counters),
            C.sizeof(A.PipelineArgs), C.sizeof(A.SynthCfg), A.EVENT24_DTYPE.itemsize]
    assert list(out[:m]) == want


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.LsgError) as e:
        pkg.Context(0)
    assert e.value.code == -5


def test_product_never_imports_the_oracle():
    prod = os.path.join(ROOT, "llm-slo-ebpf-toolkit_b200")
    for dirpath, _, files in os.walk(prod):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"(from|import)\s+oracle|oracle/|liboracle", text), f


def test_default_model_matches_oracle(lib):
    from oracle import cpu as O
    m = A.Model()
    lib.lsg_default_model(C.byref(m))
    assert bytes(m) == bytes(O.default_model())


def test_synth_host_is_deterministic_and_valid(lib):
    cfg = synth.config(5000, 500, 16, 8, 2, seed=7)
    a, b = synth.generate_host(cfg), synth.generate_host(cfg)
    for k in a:
        assert a[k].tobytes() == b[k].tobytes()
    ev, sp, ids = a["events"], a["spans"], a["idents"]
    assert ev["ident"].max() < len(ids) and sp["ident"].max() < len(ids)
    assert (sp["trace"] == np.arange(1, 501)).all()
    assert set(np.unique(ev["trace_sig"] & 15)) <= {0, 1, 2, 3, 4, 5, 7, 8, 9, 15}
    assert (ids["grp"] == np.arange(16) // 2 + 1).all()
    other = synth.generate_host(synth.config(5000, 500, 16, 8, 2, seed=8))
    assert other["events"].tobytes() != ev.tobytes()
    with pytest.raises(ValueError):
        synth.generate_host(synth.config(10, 10, 7, 8, 2))


def test_event24_pack_round_trip_and_limits(lib):
    """The 24-byte upload form (lsg_event24): packing is exact inside the narrow fields' limits, the zero Time and
    times the offset cannot express come back as the zero Time, and records that do not fit are counted."""
    cfg = synth.config(20000, 2000, 64, 8, 2, seed=3)
    ev = synth.generate_host(cfg)["events"]
    base = int(cfg.t0_ns) - 2_000_000_000
    ev["ts_ns"][:50] = A.TS_ZERO
    ev["ts_ns"][50:60] = base - 1                       # before the base
    ev["ts_ns"][60:70] = base + (1 << 42)               # beyond the offset's reach
    ev["pid"][100] = (1 << 22) - 1
    ev["conn"][101] = (1 << 20) - 1
    packed = np.zeros(len(ev), A.EVENT24_DTYPE)
    assert lib.lsg_pack_events24(ev.ctypes.data, len(ev), base, packed.ctypes.data) == 0
    back = np.zeros(len(ev), A.EVENT_DTYPE)
    assert lib.lsg_unpack_events24(packed.ctypes.data, len(ev), base, back.ctypes.data) == 0
    want = ev.copy()
    want["ts_ns"][50:70] = A.TS_ZERO
    assert back.tobytes() == want.tobytes()
    bad = ev.copy()
    bad["ident"][5] = 1 << 14
    bad["pid"][6] = 1 << 22
    bad["conn"][7] = 1 << 20
    bad["trace_sig"][8] = (1 << 26) << 4
    assert lib.lsg_pack_events24(bad.ctypes.data, len(bad), base, packed.ctypes.data) == 4
    lib.lsg_unpack_events24(packed.ctypes.data, len(bad), base, back.ctypes.data)
    assert (back["ts_ns"][5:9] == A.TS_ZERO).all() and np.array_equal(back["trace_sig"][5:9] & 15, bad["trace_sig"][5:9] & 15)
