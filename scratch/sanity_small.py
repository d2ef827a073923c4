import sys; sys.path.insert(0, '.')
import numpy as np
import llm_slo_ebpf_toolkit_b200 as pkg
from llm_slo_ebpf_toolkit_b200 import synth, _abi as A
ctx = pkg.Context(0)
# counting path, radix path (wide time range), big-span path (dense window), pairs, features, attribute, confusion, stream rows
for cfg, win in ((synth.config(30000, 3000, 8, 4, 2, seed=1, horizon_ns=10_000_000_000), 0),
                 (synth.config(30000, 3000, 8, 4, 2, seed=2, horizon_ns=4_000_000_000_000_000), 0),
                 (synth.config(30000, 300, 4, 4, 2, seed=3, horizon_ns=1_000_000_000), 5_000_000_000)):
    w = synth.generate_host(cfg)
    r = ctx.process_batch(w["spans"], w["events"], w["idents"], win, 0.7, 3)
    sig = ctx.incident_features(r["attr"], w["spans"]["incident"], len(w["inc_actual"]), want_features=True)[0]
    post, top1, elev = ctx.attribute(sig)
    c = ctx.confusion(w["inc_actual"], np.where(top1 == 255, 7, top1).astype(np.uint8), w["inc_expected"], w["inc_expected_n"], post, 0.1, 0.01)
    print("ok", int(r["n_cand"].sum()), c.n, ctx.last_sort_info())
    ctx.match_pairs(w["spans"][:300], w["events"][:300], w["idents"], 0, 0.7)
    ctx.retry_storm(w["events"], w["idents"], 10**10, 5)
    ctx.signal_status(w["events"])
ctx.decode_ringbuf(bytes(40 * 1000))
ctx.close()
print("done")
