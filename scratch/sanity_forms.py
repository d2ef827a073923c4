"""Small batches through every form of the sort + join (paths 1, 2, 3) and the radix digit widths against the CPU
oracle; seconds on a B200: python scratch/sanity_forms.py"""
import sys; sys.path.insert(0, '.')
import numpy as np
import llm_slo_ebpf_toolkit_b200 as pkg
from llm_slo_ebpf_toolkit_b200 import synth
from oracle import cpu as O
ctx = pkg.Context(0)
S = 10**9
shapes = [synth.config(40000, 2000, 8, 4, 2, seed=1, horizon_ns=20 * S),            # 20 events per span: several tiles per segment
          synth.config(30000, 3000, 64, 8, 2, seed=2, horizon_ns=3600 * S),         # sparse
          synth.config(30000, 3000, 4096, 16, 2, seed=4, horizon_ns=4000 * S),      # 29-bit keys: radix with 10-bit digits
          synth.config(30000, 300, 4, 4, 2, seed=3, horizon_ns=1 * S)]              # spans larger than a tile
for cfg in shapes:
    w = synth.generate_host(cfg)
    want = O.process_batch(w["spans"], w["events"], w["idents"], indexed=True, nthreads=4)
    for path in (2, 3, 1):
        ctx.set_join_path(path)
        got = ctx.process_batch(w["spans"], w["events"], w["idents"])
        ok = all(np.array_equal(got[k], want[k]) for k in ("cand_event", "cand_tier", "n_cand")) and got["debug"].as_tuple() == want["debug"].as_tuple()
        print("path", path, "ok" if ok else "MISMATCH", ctx.last_sort_info())
        assert ok
ctx.set_join_path(-1)
ctx.close()
print("done")
