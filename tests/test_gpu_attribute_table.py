"""Attribution of large batches reads the posteriors from a table with one row per elevated-signal mask (4096 rows, filled
by the same device function that small batches run per incident, csrc/attribute.cu).  The two forms must return the
same bits; both must equal the oracle (labels and masks exactly, posteriors to 1e-6), also after the model changes
(the table is rebuilt) and for incidents without signals (the rule path, bayesian.go:318-320)."""
import numpy as np
import pytest

from llm_slo_ebpf_toolkit_b200 import _abi as A
from oracle import cpu as O

pytestmark = pytest.mark.gpu


def _signals(n, seed):
    rng = np.random.default_rng(seed)
    sig = rng.uniform(0, 120, (n, A.NSIG))
    sig[rng.random((n, A.NSIG)) < 0.3] = np.nan
    sig[::997] = np.nan                      # incidents with no signal at all
    return sig


def test_table_form_equals_per_incident_form_and_oracle(ctx):
    n = 200_000                              # >= 65 536: the table form
    sig = _signals(n, 1)
    rule = (np.arange(n) % 8).astype(np.uint8)
    post, top1, elev = ctx.attribute(sig, rule_label=rule)
    parts = [ctx.attribute(sig[a:a + 50_000], rule_label=rule[a:a + 50_000]) for a in range(0, n, 50_000)]   # per-incident form
    assert np.array_equal(post, np.concatenate([p[0] for p in parts]), equal_nan=True)
    assert np.array_equal(top1, np.concatenate([p[1] for p in parts]))
    assert np.array_equal(elev, np.concatenate([p[2] for p in parts]))
    opost, otop1, oelev = O.attribute(O.default_model(), sig, rule_label=rule)
    assert np.array_equal(top1, otop1) and np.array_equal(elev, oelev)
    assert np.allclose(post, opost, rtol=1e-6, atol=0, equal_nan=True)
    assert np.isnan(post[::997]).all() and np.array_equal(top1[::997], rule[::997])


def test_table_follows_the_model(ctx):
    lib = ctx._lib
    import ctypes as C
    n = 100_000
    sig = _signals(n, 2)
    before = ctx.attribute(sig)
    m = A.Model()
    lib.lsg_default_model(C.byref(m))
    m.prior[0], m.prior[3] = 0.5, 0.02
    m.lik[0][0] = 0.2
    ctx._check(lib.lsg_set_model(ctx._h, C.byref(m)))
    try:
        post, top1, elev = ctx.attribute(sig)
        opost, otop1, oelev = O.attribute(m, sig)
        assert np.array_equal(top1, otop1) and np.array_equal(elev, oelev)
        assert np.allclose(post, opost, rtol=1e-6, atol=0, equal_nan=True)
        assert not np.array_equal(post, before[0], equal_nan=True)
    finally:
        d = A.Model()
        lib.lsg_default_model(C.byref(d))
        ctx._check(lib.lsg_set_model(ctx._h, C.byref(d)))
    again = ctx.attribute(sig)
    assert np.array_equal(again[0], before[0], equal_nan=True) and np.array_equal(again[1], before[1])
