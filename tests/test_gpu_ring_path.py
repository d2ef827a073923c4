"""The two forms of the permutation path's tile kernel, each forced (lsg_set_join_path(ctx, 2): sliding tile, a CTA
walks a segment of spans and keeps the events two consecutive tiles share; 3: fixed chunks of 128 spans), against the
indexed CPU oracle, field for field.  Unforced, the library picks one by the batch's events per span, so the other
suites only ever see one of them per shape.  Shapes: the common one, sparse groups with 64-bit time deltas, a window far
above the tier limits on the wide register list, one dense group whose spans exceed a tile (warp-per-span fallback), many
spans per tile, the service+node tier enumerated, narrow windows, nothing above the threshold, dead records and mixed
classes, and 20 events per span (where the sliding tile is the default)."""
import numpy as np
import pytest

from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200 import synth
from oracle import cpu as O

pytestmark = pytest.mark.gpu
S = 1_000_000_000


@pytest.fixture(params=[2, 3], ids=["sliding", "fixed"])
def forced(ctx, request):
    ctx.set_join_path(request.param)
    yield ctx
    ctx.set_join_path(-1)


def _eq(got, want):
    for k in ("cand_event", "cand_tier", "conf", "retrieval", "n_cand"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(np.nan_to_num(got["attr"], nan=-1e300), np.nan_to_num(want["attr"], nan=-1e300)), "attr"
    assert got["debug"].as_tuple() == want["debug"].as_tuple()


@pytest.mark.parametrize("n_ev,n_sp,window_ms,thr,fanout,horizon_s,n_pods,ppg", [
    (200_000, 20_000, 2000, 0.7, 3, 36, 64, 2),
    (200_000, 20_000, 2000, 0.7, 3, 3600, 4096, 2),
    (35_000, 1_800, 120_000, 0.8, 8, 3600, 2, 2),
    (400_000, 4_000, 2000, 0.7, 3, 20, 2, 2),
    (300_000, 60_000, 2000, 0.7, 3, 30, 8, 2),
    (200_000, 20_000, 2000, 0.6, 12, 36, 16, 2),
    (200_000, 20_000, 50, 0.7, 3, 20, 16, 2), (200_000, 20_000, 400, 0.95, 2, 20, 16, 1),
    (100_000, 10_000, 2000, 1.01, 3, 36, 16, 4),
    (600_000, 30_000, 2000, 0.7, 3, 36, 64, 2),       # 20 events per span
    (1_000, 3, 2000, 0.7, 3, 36, 4, 2), (5, 700, 2000, 0.7, 3, 36, 4, 2),   # fewer spans than a chunk, fewer events than spans
])
def test_forced_tile_form_equals_the_oracle(forced, n_ev, n_sp, window_ms, thr, fanout, horizon_s, n_pods, ppg):
    cfg = synth.config(n_ev, n_sp, n_pods, max(ppg, 8 if n_pods >= 8 else n_pods), ppg, seed=window_ms + fanout,
                       horizon_ns=horizon_s * S, multi_fault=fanout % 2)
    w = synth.generate_host(cfg)
    got = forced.process_batch(w["spans"], w["events"], w["idents"], window_ms * 1_000_000, thr, fanout)
    want = O.process_batch(w["spans"], w["events"], w["idents"], window_ms * 1_000_000, thr, fanout, indexed=True, nthreads=8)
    _eq(got, want)


def test_forced_tile_form_dead_records_and_mixed_classes(forced):
    cfg = synth.config(150_000, 15_000, 16, 8, 2, seed=5, horizon_ns=30 * S)
    w = synth.generate_host(cfg)
    sp, ev, ids = w["spans"].copy(), w["events"].copy(), w["idents"].copy()
    rng = np.random.default_rng(9)
    sp["ts_ns"][rng.choice(len(sp), 700, replace=False)] = A.TS_ZERO
    ev["ts_ns"][rng.choice(len(ev), 3000, replace=False)] = A.TS_ZERO
    ev["ts_ns"][rng.choice(len(ev), 2000, replace=False)] += 400 * 86400 * S
    ev["trace_sig"][rng.choice(len(ev), 1000, replace=False)] |= 15
    ids["grp"][3] = 0
    ids["service"][::2] = 0
    got = forced.process_batch(sp, ev, ids)
    want = O.process_batch(sp, ev, ids, indexed=True, nthreads=8)
    _eq(got, want)
    empty = forced.process_batch(sp, ev[:0], ids)
    assert (empty["cand_event"] == A.NONE).all() and empty["debug"].as_tuple() == (0, 0, 0, 0)
    dead = sp.copy()
    dead["ts_ns"][:] = A.TS_ZERO
    got = forced.process_batch(dead, ev, ids)
    want = O.process_batch(dead, ev, ids, indexed=True, nthreads=8)
    _eq(got, want)


def test_the_forms_agree_on_a_large_radix_sorted_batch(ctx):
    """C3's shape at 1/50 of its size (radix path, 4096 pods, 20 events per span): both forms, identical outputs."""
    cfg = synth.config(20_000_000, 1_000_000, 4096, 16, 2, seed=3, horizon_ns=360 * S, multi_fault=1)
    w = synth.generate_host(cfg)
    out = {}
    try:
        for path in (2, 3):
            ctx.set_join_path(path)
            out[path] = ctx.process_batch(w["spans"], w["events"], w["idents"])
    finally:
        ctx.set_join_path(-1)
    for k in ("cand_event", "cand_tier", "conf", "retrieval", "n_cand"):
        assert np.array_equal(out[2][k], out[3][k]), k
    assert np.array_equal(out[2]["attr"], out[3]["attr"], equal_nan=True)
    assert out[2]["debug"].as_tuple() == out[3]["debug"].as_tuple() and out[2]["debug"].as_tuple()[0] > 1_000_000
