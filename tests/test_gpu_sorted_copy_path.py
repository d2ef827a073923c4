"""The comparison path kept behind LSG_JOIN_SORTED_COPY=1 (column gather + TMA-staged join) must give the same results
as the default path (the join gathers its own tiles).  The switch is read once per process, so it runs in a child."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys
sys.path.insert(0, %r)
import numpy as np
import llm_slo_ebpf_toolkit_b200 as pkg
from llm_slo_ebpf_toolkit_b200 import synth
from oracle import cpu as O
ctx = pkg.Context(0)
for cfg, win, thr, fan in ((synth.config(120000, 12000, 16, 8, 2, seed=11, horizon_ns=30_000_000_000), 0, 0.7, 3),
                           (synth.config(60000, 6000, 8, 4, 2, seed=12, horizon_ns=4_000_000_000_000_000), 0, 0.6, 8),
                           (synth.config(30000, 300, 4, 4, 2, seed=13, horizon_ns=1_000_000_000), 5_000_000_000, 0.7, 12)):
    w = synth.generate_host(cfg)
    got = ctx.process_batch(w["spans"], w["events"], w["idents"], win, thr, fan)
    want = O.process_batch(w["spans"], w["events"], w["idents"], win, thr, fan, indexed=True, nthreads=8)
    for k in ("cand_event", "cand_tier", "conf", "retrieval", "n_cand"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(np.nan_to_num(got["attr"], nan=-1e300), np.nan_to_num(want["attr"], nan=-1e300))
    assert got["debug"].as_tuple() == want["debug"].as_tuple()
print("same")
""" % ROOT


@pytest.mark.gpu
@pytest.mark.parametrize("sorted_copy", ["1", ""])
def test_both_join_paths_match_the_oracle(sorted_copy):
    env = dict(os.environ)
    env.pop("LSG_JOIN_SORTED_COPY", None)
    if sorted_copy:
        env["LSG_JOIN_SORTED_COPY"] = sorted_copy
    p = subprocess.run([sys.executable, "-c", CHILD], capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert p.returncode == 0 and "same" in p.stdout, p.stdout + p.stderr
