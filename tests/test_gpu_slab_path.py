"""The second implementation of the sort + join stages (lsg_set_join_path(ctx, 1): bucket partition of the records,
then the slab join on contiguous tiles with an in-tile counting sort by time quantum) against the indexed CPU oracle,
field for field, on the shapes that reach its special paths: 32-bit and 64-bit time deltas, quanta wider than the
inner window, several sub-chunks of spans per tile, buckets larger than a tile (warp-per-span fallback), note-list
overflow, fanout above the register list, spans and events that can match nothing, the two-level partition."""
import numpy as np
import pytest

from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200 import synth
from oracle import cpu as O

pytestmark = pytest.mark.gpu
S = 1_000_000_000


@pytest.fixture()
def slab(ctx):
    ctx.set_join_path(1)
    yield ctx
    ctx.set_join_path(-1)


def _eq(got, want):
    for k in ("cand_event", "cand_tier", "conf", "retrieval", "n_cand"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(np.nan_to_num(got["attr"], nan=-1e300), np.nan_to_num(want["attr"], nan=-1e300)), "attr"
    assert got["debug"].as_tuple() == want["debug"].as_tuple()


@pytest.mark.parametrize("n_ev,n_sp,window_ms,thr,fanout,horizon_s,n_pods,ppg", [
    (200_000, 20_000, 2000, 0.7, 3, 36, 64, 2),       # the common shape
    (200_000, 20_000, 2000, 0.7, 3, 3600, 4096, 2),   # sparse: coarsened quantum, 64-bit deltas, one bucket per group
    (35_000, 1_800, 120_000, 0.8, 8, 3600, 2, 2),     # window far above the tier limits, fanout on the wide register list
    (400_000, 4_000, 2000, 0.7, 3, 20, 2, 2),         # dense: 10 k events per second in one group -> buckets above a tile
    (300_000, 60_000, 2000, 0.7, 3, 30, 8, 2),        # many spans per tile: several sub-chunks
    (200_000, 20_000, 2000, 0.6, 12, 36, 16, 2),      # service+node tier enumerated, fanout above the register list
    (200_000, 20_000, 50, 0.7, 3, 20, 16, 2), (200_000, 20_000, 400, 0.95, 2, 20, 16, 1),
    (100_000, 10_000, 2000, 1.01, 3, 36, 16, 4),      # nothing reaches the threshold
])
def test_slab_path_equals_the_oracle(slab, n_ev, n_sp, window_ms, thr, fanout, horizon_s, n_pods, ppg):
    cfg = synth.config(n_ev, n_sp, n_pods, max(ppg, 8 if n_pods >= 8 else n_pods), ppg, seed=window_ms + fanout,
                       horizon_ns=horizon_s * S, multi_fault=fanout % 2)
    w = synth.generate_host(cfg)
    got = slab.process_batch(w["spans"], w["events"], w["idents"], window_ms * 1_000_000, thr, fanout)
    want = O.process_batch(w["spans"], w["events"], w["idents"], window_ms * 1_000_000, thr, fanout, indexed=True, nthreads=8)
    _eq(got, want)


def test_slab_path_dead_records_and_mixed_classes(slab):
    cfg = synth.config(150_000, 15_000, 16, 8, 2, seed=5, horizon_ns=30 * S)
    w = synth.generate_host(cfg)
    sp, ev, ids = w["spans"].copy(), w["events"].copy(), w["idents"].copy()
    rng = np.random.default_rng(9)
    sp["ts_ns"][rng.choice(len(sp), 700, replace=False)] = A.TS_ZERO          # spans with the zero Time
    ev["ts_ns"][rng.choice(len(ev), 3000, replace=False)] = A.TS_ZERO
    ev["ts_ns"][rng.choice(len(ev), 2000, replace=False)] += 400 * 86400 * S   # far outside the spans' range
    ev["trace_sig"][rng.choice(len(ev), 1000, replace=False)] |= 15            # unsupported signal type
    ids["grp"][3] = 0                                                          # a row that matches nothing
    ids["service"][::2] = 0                                                    # mixed classes inside every group
    got = slab.process_batch(sp, ev, ids)
    want = O.process_batch(sp, ev, ids, indexed=True, nthreads=8)
    _eq(got, want)
    # no events at all, and no span that can match
    empty = slab.process_batch(sp, ev[:0], ids)
    assert (empty["cand_event"] == A.NONE).all() and empty["debug"].as_tuple() == (0, 0, 0, 0)
    dead = sp.copy()
    dead["ts_ns"][:] = A.TS_ZERO
    got = slab.process_batch(dead, ev, ids)
    want = O.process_batch(dead, ev, ids, indexed=True, nthreads=8)
    _eq(got, want)


def test_slab_path_two_level_partition(slab):
    """More than 2^17 buckets: the records are scattered twice (super-buckets, then buckets)."""
    cfg = synth.config(3_000_000, 150_000, 4096, 16, 2, seed=11, horizon_ns=360 * S, multi_fault=1)
    w = synth.generate_host(cfg)
    got = slab.process_batch(w["spans"], w["events"], w["idents"])
    assert slab.last_sort_info()["key_bits"] >= 23
    want = O.process_batch(w["spans"], w["events"], w["idents"], indexed=True, nthreads=8)
    _eq(got, want)
