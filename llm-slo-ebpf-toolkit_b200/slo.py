"""Quantiles of the SLO calculator on the GPU.

Mirrors `quantile` / `Aggregate` of pkg/slo/calculator.go:99-149 and the node p95 of
pkg/releasegate/gate.go:345-353: same names, same argument meaning, same results (0 for an empty list, q clamped
to [0, 1], linear interpolation between the two neighbours of q * (n - 1)).  The values are sorted by the library's
radix sort; there is no CPU path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Iterable, Sequence

import numpy as np

from ._lib import Context


def quantile(ctx: Context, values: Sequence[float], q: float) -> float:
    """calculator.go:121-149"""
    out, _ = ctx.segment_quantiles(np.asarray(values, np.float64), None, 1, [q])
    return float(out[0, 0])


def segment_quantiles(ctx: Context, values, segment, n_segments: int, qs: Iterable[float]):
    """quantile() of every list at once: record i belongs to list segment[i]; returns (out[n_segments, len(qs)], counts)"""
    return ctx.segment_quantiles(values, segment, n_segments, list(qs))


@dataclass
class Percentiles:
    """calculator.go:32-40"""
    TTFTP50: float = 0.0
    TTFTP95: float = 0.0
    TTFTP99: float = 0.0
    TokensPerSP50: float = 0.0
    TokensPerSP95: float = 0.0
    RetrievalP95MS: float = 0.0


def Aggregate(ctx: Context, ttft_ms, tokens_per_s, retrieval_ms) -> Percentiles:
    """calculator.go:99-118: negative TTFT / tokens-per-second count as 0 (nonNegative, calculator.go:151-156);
    `retrieval_ms` is TotalRetrievalMS per item."""
    ttft = np.asarray(ttft_ms, np.float64)
    tps = np.asarray(tokens_per_s, np.float64)
    ret = np.asarray(retrieval_ms, np.float64)
    if not (len(ttft) == len(tps) == len(ret)):
        raise ValueError("one value of each metric per item")
    ttft = np.where(ttft < 0, 0.0, ttft)
    tps = np.where(tps < 0, 0.0, tps)
    n = len(ttft)
    values = np.concatenate([ttft, tps, ret])
    segment = np.repeat(np.arange(3, dtype=np.uint32), n)
    out, _ = ctx.segment_quantiles(values, segment, 3, [0.50, 0.95, 0.99])
    return Percentiles(TTFTP50=out[0, 0], TTFTP95=out[0, 1], TTFTP99=out[0, 2], TokensPerSP50=out[1, 0],
                       TokensPerSP95=out[1, 1], RetrievalP95MS=out[2, 1])


def node_p95(ctx: Context, node_of_sample, overhead_pct, n_nodes: int) -> Dict[int, float]:
    """releasegate/gate.go:345-353: p95 of the observed overhead per node (nodes without samples are left out)"""
    out, cnt = ctx.segment_quantiles(overhead_pct, node_of_sample, n_nodes, [0.95])
    return {int(k): float(out[k, 0]) for k in range(n_nodes) if cnt[k]}
