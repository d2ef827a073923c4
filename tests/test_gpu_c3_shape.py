"""BASELINE.json configs[2] (C3: 1 B events x 50 M spans, 4 096 pods, multi-fault incidents) at its shape and a
tenth of its size — 100 M events x 5 M spans, the largest the indexed CPU oracle finishes in about a minute — through
the device-resident pipeline, compared exactly: kept candidates and tiers of every span, the debug counters, the
per-incident maxima, labels and the confusion block; posteriors to 1e-6.  This is the configuration that takes the
radix path of the permutation sort (29-bit keys: more key values than the counting sort's table) and the two-level
scatter of the bucket partition, so both implementations of the sort + join are run."""
import ctypes as C

import numpy as np
import pytest
import torch

from llm_slo_ebpf_toolkit_b200 import _abi as A
from llm_slo_ebpf_toolkit_b200 import synth
from llm_slo_ebpf_toolkit_b200.device import DeviceWorkload
from oracle import cpu as O

pytestmark = pytest.mark.gpu
SECOND = 1_000_000_000


@pytest.fixture(scope="module")
def c3s_oracle():
    cfg = synth.config(100_000_000, 5_000_000, 4096, 16, 2, 10, 1, horizon_ns=360 * SECOND)
    w = O.generate_host(cfg)
    threads = max(1, O.hw_threads())
    j = O.process_batch(w["spans"], w["events"], w["idents"], indexed=True, nthreads=threads)
    ni = len(w["inc_actual"])
    sig = O.incident_features(j["attr"], w["spans"]["incident"], ni)
    post, top1, _ = O.attribute(O.default_model(), sig)
    conf = O.confusion(w["inc_actual"], top1, w["inc_expected"], w["inc_expected_n"], post, 0.10, 0.01)
    want = {"cand_event": j["cand_event"], "cand_tier": j["cand_tier"], "debug": j["debug"].as_tuple(), "sig": sig,
            "post": post, "top1": top1, "conf": bytes(conf), "n_inc": ni}
    del j, w
    return cfg, want


@pytest.mark.parametrize("path", [0, 1])
def test_c3_shape_exact(ctx, c3s_oracle, path):
    cfg, want = c3s_oracle
    ctx.set_join_path(path)
    try:
        wl = DeviceWorkload(ctx, cfg)          # generated on the device by the same counter-based generator
        wl.step()
        torch.cuda.synchronize()
        c = wl.read_counters()
        info = ctx.last_sort_info()
        assert info["key_bits"] > 25, info     # beyond the counting sort's table: radix keys / two scatter levels
        assert c.bad_idents == 0
        f = 3
        got_e = wl.host("cand_event", np.uint32)[: wl.n_spans * f].reshape(-1, f)
        got_t = wl.host("cand_tier", np.uint8)[: wl.n_spans * f].reshape(-1, f)
        assert np.array_equal(got_e, want["cand_event"])
        assert np.array_equal(got_t, want["cand_tier"])
        assert c.debug.as_tuple() == want["debug"]
        sig = wl.host("inc_signals", np.float64)[: want["n_inc"] * A.NSIG].reshape(-1, A.NSIG)
        assert np.array_equal(sig, want["sig"], equal_nan=True)
        assert np.array_equal(wl.host("top1", np.uint8)[: want["n_inc"]], want["top1"])
        post = wl.host("posterior", np.float64)[: want["n_inc"] * A.NDOM].reshape(-1, A.NDOM)
        assert np.allclose(post, want["post"], rtol=1e-6, atol=0, equal_nan=True)
        assert bytes(c.confusion) == want["conf"]
        del wl
        torch.cuda.empty_cache()
    finally:
        ctx.set_join_path(-1)
