"""The `assertions:` blocks of the incident-lab scenarios (test/incident-lab/scenarios/<name>.yaml, e.g.
dns_latency.yaml:34-46) evaluated on batches of signal values (SURVEY.md §8f, F4's second half).

An assertion is {metric, operator, value, phase}.  Metrics named `<series>_p50` / `_p95` / `_p99` are quantiles of
the series' values in that phase: all of them, for every phase and series of a run, come out of ONE call of
lsg_segment_quantiles (quantile() of pkg/slo/calculator.go:121-149, one list per (phase, series)).  Any other metric
name (`error_rate`, `correlation_enrichment_rate`, `attribution_recall`, a bare signal name such as `cpu_steal_pct`,
...) is looked up in `scalars[phase]`; a bare series name falls back to the mean of the series, which is what the
scenario's fault_profile states for it.  The reference holds these blocks as data only (no Go code reads them), so
the evaluator's semantics beyond quantile() are this library's reading of the YAML: parity unpinned for them.
"""
from __future__ import annotations

import operator as _op
import re
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np

from ._lib import Context

_OPS = {">": _op.gt, ">=": _op.ge, "<": _op.lt, "<=": _op.le, "==": _op.eq, "!=": _op.ne}
_Q = re.compile(r"^(?P<series>.+)_p(?P<q>50|90|95|99)$")


@dataclass
class AssertionResult:
    metric: str
    operator: str
    value: float
    phase: str
    observed: Optional[float]   # None: the run holds neither the series nor a scalar of that name
    passed: Optional[bool]


def load_assertions(doc) -> List[dict]:
    """`doc`: the parsed YAML of a scenario (or its `assertions` list)."""
    items = doc.get("assertions", []) if isinstance(doc, dict) else doc
    out = []
    for a in items or []:
        if a.get("operator") not in _OPS:
            raise ValueError("unknown operator %r in assertion on %r" % (a.get("operator"), a.get("metric")))
        out.append({"metric": str(a["metric"]), "operator": a["operator"], "value": float(a["value"]),
                    "phase": str(a.get("phase") or "fault")})
    return out


def evaluate(ctx: Context, assertions: Sequence[dict], series: Dict[str, Dict[str, Sequence[float]]],
             scalars: Optional[Dict[str, Dict[str, float]]] = None) -> List[AssertionResult]:
    """series[phase][name] = the values of one series in one phase; scalars[phase][metric] = already reduced metrics."""
    scalars = scalars or {}
    wanted = {}   # (phase, series) -> list id
    for a in assertions:
        m = _Q.match(a["metric"])
        if m and m.group("series") in series.get(a["phase"], {}):
            wanted.setdefault((a["phase"], m.group("series")), len(wanted))
    quant = None
    if wanted:
        vals = [np.asarray(series[ph][name], np.float64) for (ph, name) in wanted]
        seg = np.concatenate([np.full(len(v), i, np.uint32) for i, v in enumerate(vals)]) if vals else np.zeros(0, np.uint32)
        quant, _ = ctx.segment_quantiles(np.concatenate(vals), seg, len(vals), [0.50, 0.90, 0.95, 0.99])
    col = {"50": 0, "90": 1, "95": 2, "99": 3}
    out = []
    for a in assertions:
        ph, metric = a["phase"], a["metric"]
        observed = None
        m = _Q.match(metric)
        if m and (ph, m.group("series")) in wanted:
            observed = float(quant[wanted[(ph, m.group("series"))], col[m.group("q")]])
        elif metric in scalars.get(ph, {}):
            observed = float(scalars[ph][metric])
        elif metric in series.get(ph, {}) and len(series[ph][metric]):
            observed = float(np.mean(np.asarray(series[ph][metric], np.float64)))
        passed = None if observed is None else bool(_OPS[a["operator"]](observed, a["value"]))
        out.append(AssertionResult(metric, a["operator"], a["value"], ph, observed, passed))
    return out
