"""F4: quantile() of pkg/slo/calculator.go:121-149 — the oracle against the reference's own test values
(calculator_test.go:34-52) and its line-faithful restatement; the GPU path against the oracle, bit for bit."""
import math

import numpy as np
import pytest

from oracle import cpu as orc
from oracle import ref_str as R


def test_oracle_matches_reference_test_values():
    # TestAggregate, calculator_test.go:34-52: three snapshots, p50 of TTFT = 200, p50 of tokens/s = 30, p95 of retrieval > 0
    ttft, tps = [100.0, 200.0, 300.0], [50.0, 30.0, 10.0]
    retrieval = [20 + 10 + 5.0, 40 + 15 + 10.0, 60 + 25 + 15.0]
    assert R.quantile(ttft, 0.50) == 200
    assert R.quantile(tps, 0.50) == 30
    assert R.quantile(retrieval, 0.95) > 0
    vals = np.array(ttft + tps + retrieval)
    seg = np.repeat(np.arange(3, dtype=np.uint32), 3)
    out, cnt = orc.segment_quantiles(vals, seg, 3, [0.5, 0.95])
    assert out[0, 0] == 200 and out[1, 0] == 30 and out[2, 1] > 0
    assert list(cnt) == [3, 3, 3]
    # interpolation by hand: pos = 0.95 * 2 = 1.9 -> 65 * 0.1 + 100 * 0.9 in the reference's operation order
    frac = 0.95 * 2.0 - 1.0
    assert out[2, 1] == 65.0 * (1 - frac) + 100.0 * frac


def test_oracle_c_equals_line_faithful_restatement():
    rng = np.random.default_rng(5)
    n, n_seg = 4000, 37
    vals = rng.lognormal(3.0, 1.0, n)
    vals[rng.integers(0, n, 40)] = np.nan
    vals[rng.integers(0, n, 10)] = np.inf
    vals[rng.integers(0, n, 10)] = -np.inf
    vals[rng.integers(0, n, 200)] = 7.25          # ties
    seg = rng.integers(0, n_seg + 3, n).astype(np.uint32)   # ids >= n_seg are ignored
    seg[seg == 5] = 6                                       # an empty list
    qs = [-0.5, 0.0, 0.25, 0.5, 0.9, 0.95, 0.99, 1.0, 1.5]
    out, cnt = orc.segment_quantiles(vals, seg, n_seg, qs)
    for s in range(n_seg):
        lst = [float(v) for v in vals[seg == s]]
        assert cnt[s] == len(lst)
        for j, q in enumerate(qs):
            want = R.quantile(lst, q)
            got = out[s, j]
            assert (math.isnan(want) and math.isnan(got)) or want == got, (s, q, want, got)
    assert cnt[5] == 0 and not out[5].any()


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return np.array_equal(a, b, equal_nan=True)


@pytest.mark.gpu
def test_gpu_quantiles_bit_exact(ctx):
    rng = np.random.default_rng(11)
    qs = [-1.0, 0.0, 0.01, 0.25, 0.5, 0.75, 0.9, 0.95, 0.99, 0.999, 1.0, 2.0]
    for n, n_seg in ((0, 4), (1, 1), (2, 1), (1000, 1), (5000, 12), (200_000, 4096), (300_000, 5000), (2_000_000, 63)):
        vals = rng.lognormal(2.0, 1.5, n) * rng.choice([1.0, -1.0], n, p=[0.9, 0.1])
        if n > 100:
            vals[rng.integers(0, n, n // 100)] = np.nan
            vals[rng.integers(0, n, 5)] = np.inf
            vals[rng.integers(0, n, 5)] = -np.inf
            vals[rng.integers(0, n, n // 20)] = 3.5
            vals[rng.integers(0, n, 5)] = 0.0
        seg = rng.integers(0, n_seg + 2, n).astype(np.uint32)
        if n_seg > 3:
            seg[seg == 2] = 3
        want, want_cnt = orc.segment_quantiles(vals, seg, n_seg, qs)
        got, got_cnt = ctx.segment_quantiles(vals, seg, n_seg, qs)
        assert np.array_equal(want_cnt, got_cnt), (n, n_seg)
        assert _same(want, got), (n, n_seg)
        if n_seg == 1:   # no segment array: one list holding every value
            want1, _ = orc.segment_quantiles(vals, None, 1, qs)
            got1, cnt1 = ctx.segment_quantiles(vals, None, 1, qs)
            assert _same(want1, got1) and cnt1[0] == n


@pytest.mark.gpu
def test_gpu_aggregate_and_errors(ctx):
    import importlib
    slo = importlib.import_module("llm_slo_ebpf_toolkit_b200.slo")
    # TestAggregate, calculator_test.go:34-52
    p = slo.Aggregate(ctx, [100, 200, 300], [50, 30, 10], [35, 65, 100])
    assert p.TTFTP50 == 200 and p.TokensPerSP50 == 30 and p.RetrievalP95MS > 0
    assert slo.quantile(ctx, [], 0.95) == 0.0          # calculator.go:122-124
    assert slo.quantile(ctx, [42.0], 0.3) == 42.0      # calculator.go:136-138
    assert slo.Aggregate(ctx, [-5.0], [-1.0], [3.0]).TTFTP50 == 0.0   # nonNegative, calculator.go:151-156
    nodes = slo.node_p95(ctx, np.array([0, 0, 2, 2, 2], np.uint32), [1.0, 3.0, 2.0, 4.0, 6.0], 3)
    assert set(nodes) == {0, 2} and nodes[0] == R.quantile([1.0, 3.0], 0.95) and nodes[2] == R.quantile([2.0, 4.0, 6.0], 0.95)
    with pytest.raises(Exception):
        ctx.segment_quantiles([1.0, 2.0], None, 1, [float("nan")])
    with pytest.raises(Exception):
        ctx.segment_quantiles([1.0, 2.0], None, 2, [0.5])
