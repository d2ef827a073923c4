"""lsg_validate_groups: the join-group contract (include/llmslo_gpu.h) checked on a batch.  The interner's groups
pass; a group split along each of the three identities Match() can accept a pair on (dns.go:58-73) is reported, and a
batch with such a split really does lose matches against the oracle, which is why the check exists."""
import numpy as np
import pytest

from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200 import synth
from llm_slo_ebpf_toolkit_b200.interner import Interner
from tests.randgen import random_batch

pytestmark = pytest.mark.gpu


def test_validate_groups(ctx):
    cfg = synth.config(100_000, 10_000, 16, 8, 2, seed=9, horizon_ns=30_000_000_000)
    w = synth.generate_host(cfg)
    assert ctx.validate_groups(w["spans"], w["events"], w["idents"]) == (0, 0, 0, 0)
    # (service, node) split: the two pods of one class get different groups
    ids = w["idents"].copy()
    ids["grp"][1] = 99
    pod, cls, tr, bad = ctx.validate_groups(w["spans"], w["events"], ids)
    assert cls >= 1 and pod == 0 and bad == 0
    # pod split: a second identity row with the same pod under another group
    ids = np.concatenate([w["idents"], w["idents"][:1]])
    ids["grp"][-1] = 77
    assert ctx.validate_groups(w["spans"], w["events"], ids)[0] == 1
    # trace split: an event that carries a span's trace id sits in another group
    ev = w["events"].copy()
    k = int(np.nonzero((ev["trace_sig"] >> 4) != 0)[0][0])
    ev["ident"][k] = (int(ev["ident"][k]) + 2) % len(w["idents"])
    assert ctx.validate_groups(w["spans"], ev, w["idents"])[2] >= 1
    # an ident outside the table; unsupported or untimed events are not looked at
    ev = w["events"].copy()
    ev["ident"][5] = 10_000
    assert ctx.validate_groups(w["spans"], ev, w["idents"])[3] == 1
    ev = w["events"].copy()
    ev["ident"][k] = (int(ev["ident"][k]) + 2) % len(w["idents"])
    ev["ts_ns"][k] = A.TS_ZERO
    assert ctx.validate_groups(w["spans"], ev, w["idents"]) == (0, 0, 0, 0)
    assert ctx.validate_groups(w["spans"][:0], w["events"][:0], w["idents"][:0]) == (0, 0, 0, 0)


def test_interner_groups_satisfy_the_contract(ctx):
    """The union-find groups of the interner (rows sharing a pod, a (service, node) pair or a trace id are merged) pass on
    random string batches, including signals that carry only a trace id."""
    from llm_slo_ebpf_toolkit_b200 import correlation as CO
    for seed in range(4):
        spans, signals = random_batch(300 + seed, 60, 400, n_pods=3 + seed, dense=seed % 2 == 0)
        it = Interner()
        sp = it.span_records([CO.SpanRef(s.TraceID, s.Service, s.Node, s.Pod, s.PID, s.ConnTuple, s.Timestamp) for s in spans])
        ev = it.event_records([CO.SignalRef(g.Signal, g.TraceID, g.Service, g.Node, g.Pod, g.PID, g.ConnTuple, g.Timestamp, g.Value)
                               for g in signals])
        assert ctx.validate_groups(sp, ev, it.ident_table()) == (0, 0, 0, 0)
