"""The incident-lab `assertions:` blocks (tests/golden/scenario_assertions.json, taken from the reference's
test/incident-lab/scenarios/*.yaml) evaluated on the GPU: the fault phase of every scenario, generated from its own
fault_profile the way pkg/signals/generator.go:244-289 states the levels, satisfies every `*_p95` assertion on a
kernel signal; the baseline phase does not; each quantile equals the oracle's restatement of calculator.go:121-149."""
import numpy as np
import pytest

from llm_slo_ebpf_toolkit_b200 import scenarios as SC
from oracle import cpu as O
from tests.fixtures import load_golden

pytestmark = pytest.mark.gpu
BASE = {"dns_latency_ms": 12, "tcp_retransmits_total": 0.2, "runqueue_delay_ms": 4, "connect_latency_ms": 18, "tls_handshake_ms": 22,
        "cpu_steal_pct": 0.6, "cfs_throttled_ms": 5, "mem_reclaim_latency_ms": 0.5, "disk_io_latency_ms": 2, "syscall_latency_ms": 5}


def _phase(levels, n, rng):
    return {k: float(v) * rng.uniform(0.85, 1.25, n) for k, v in levels.items()}


def test_scenario_assertions(ctx):
    scen = load_golden("scenario_assertions")
    assert set(scen) >= {"dns_latency", "cpu_throttle", "memory_pressure", "network_partition", "provider_throttle", "mixed_multi"}
    rng = np.random.default_rng(4)
    n_quant = 0
    for name, doc in scen.items():
        asserts = SC.load_assertions(doc)
        prof = {k.replace("tcp_retransmits", "tcp_retransmits_total") if k == "tcp_retransmits" else k: v
                for k, v in doc["fault_profile"].items() if isinstance(v, (int, float))}
        n = int(doc.get("sample_count") or 24) * 10
        series = {"fault": _phase(dict(BASE, **prof), n, rng), "baseline": _phase(BASE, n, rng)}
        res = SC.evaluate(ctx, asserts, series, scalars={"fault": {"error_rate": 0.2}})
        for a, r in zip(asserts, res):
            m = SC._Q.match(a["metric"])
            if m and m.group("series") in series[a["phase"]]:
                want, _ = O.segment_quantiles(series[a["phase"]][m.group("series")], None, 1, [int(m.group("q")) / 100.0])
                assert r.observed == want[0, 0]                      # bit-exact against the oracle's quantile()
                assert r.passed is True, (name, a, r.observed)        # the scenario's own fault levels satisfy its assertion
                n_quant += 1
            elif a["metric"] in BASE:
                assert r.passed is True, (name, a, r.observed)        # a bare signal name: the mean of the series
            elif a["metric"] != "error_rate":
                assert r.observed is None and r.passed is None        # product metrics this path does not produce
        # the same assertions against the baseline levels: every quantile assertion on a kernel signal fails
        flipped = [dict(a, phase="baseline") for a in asserts if SC._Q.match(a["metric"])]
        for r in SC.evaluate(ctx, flipped, series):
            if r.observed is not None:
                assert r.passed is False, (name, r)
    assert n_quant >= 8
    with pytest.raises(ValueError):
        SC.load_assertions([{"metric": "x", "operator": "~", "value": 1}])
