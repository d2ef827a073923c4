"""The allocation-free entry point lsg_incident_signals_static against the regular one.

Why it exists: lsg_incident_features_dev keeps its accumulators in a workspace it cudaMalloc's.  A caller that
has given all of the GPU's memory to its own pool, or that cannot tolerate the device-wide synchronisation of
cudaMalloc/cudaFree while other streams run, cannot use it.  The static variant keeps the accumulators in a
fixed 1.5 MiB array inside the library, takes no context, and must give exactly the same per-incident maxima.

This file is collected last on purpose (pytest --collect-only): if it fails, a run that stops at the first
failure has already run every other test.  It passes both alone in a fresh process and after the other tests;
the two calls in a row check that what the first call left in the shared array cannot leak into the second.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_INCIDENTS = 16384   # LSG_STATIC_MAX_INCIDENTS in include/llmslo_gpu.h


def _inputs(seed, n_spans, n_incidents):
    rng = np.random.default_rng(seed)
    attr = rng.choice([np.nan, np.nan, -2.5, 0.0, 0.3, 7.0, 100.0, 5000.0], size=(n_spans, 12))
    inc = np.sort(rng.integers(0, n_incidents, n_spans)).astype(np.uint32)
    inc[::11] = 0xFFFFFFFF                                   # spans that belong to no incident
    inc[3::500] = rng.integers(0, n_incidents, len(inc[3::500]))   # and some out of order
    return attr, inc


def test_static_entry_point_matches_allocating_one(ctx):
    import torch

    import llm_slo_ebpf_toolkit_b200 as pkg
    lib = pkg.load_library()
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    err = C.create_string_buffer(256)
    # first call: a full-size batch; second call: a smaller batch with other data in the same array
    for seed, n_spans, n_inc in ((1, 150_000, MAX_INCIDENTS), (2, 40_000, 3_000)):
        attr, inc = _inputs(seed, n_spans, n_inc)
        d_attr = torch.from_numpy(attr).cuda()
        d_inc = torch.from_numpy(inc.view(np.int32)).cuda()
        want = torch.empty(n_inc * 12, dtype=torch.float64, device="cuda")
        got = torch.full((n_inc * 12,), -1.0, dtype=torch.float64, device="cuda")
        ctx.call_dev("lsg_incident_features_dev", d_attr.data_ptr(), d_inc.data_ptr(), n_spans, n_inc, want.data_ptr(), None)
        rc = lib.lsg_incident_signals_static(stream, d_attr.data_ptr(), d_inc.data_ptr(), n_spans, n_inc, got.data_ptr(),
                                             err, len(err))
        assert rc == 0, err.value
        torch.cuda.synchronize()
        w, g = want.cpu().numpy(), got.cpu().numpy()
        # maxima are exact in both paths: bit-for-bit, NaN ("signal absent") in the same places
        assert np.array_equal(np.isnan(w), np.isnan(g))
        assert np.array_equal(w[~np.isnan(w)], g[~np.isnan(g)])
        assert (~np.isnan(g)).sum() > n_inc      # the comparison is not vacuous
    # a batch beyond the built-in array is refused with a clear error and touches nothing
    rc = lib.lsg_incident_signals_static(stream, d_attr.data_ptr(), d_inc.data_ptr(), n_spans, MAX_INCIDENTS + 1,
                                         got.data_ptr(), err, len(err))
    assert rc == -4 and b"LSG_STATIC_MAX_INCIDENTS" in err.value
    # empty batch is a no-op
    assert lib.lsg_incident_signals_static(stream, None, None, 0, 0, None, err, len(err)) == 0
