#!/usr/bin/env python3
"""bench.py — correlated + attributed kernel events per second on N x B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's algorithm on the host cores

A step = one pass of the whole hot path over one batch of synthetic input:
sort events -> sort spans -> windowed join (Match, top-3, per-attribute max, counters) -> per-incident
features -> Bayesian attribution -> confusion matrix [-> sum-allreduce of the counter block when N > 1].
Workload at N = 1: BASELINE.json configs[1] ("C2": 10 M kernel events x 1 M spans, 64 pods, 9 signal types,
single-fault incidents); at N > 1 every rank holds its own C2-sized shard of distinct pods (weak scaling,
the partition key is the pod/node identity so the join is shard-local).

`value`      events/s with the inputs resident in HBM (device events around K steps, max over ranks)
`e2e`        the same metric through lsg_pipeline with HOST (pinned) buffers: H2D of the records and D2H of
             every per-span / per-incident result inside the timed region
`roofline`   the dominant kernel's algorithmic bytes per launch / its mean duration (CUDA events recorded on
             the launch stream inside the timed region) against MEASURED_PEAKS.json
`cpu_baseline` the CPU oracle (indexed variant, all host threads) on a bounded sample of the same workload
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "correlated_and_attributed_events_per_sec"
UNIT = "events/s"
SECOND = 1_000_000_000


def measured_peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "50"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, reasons = [], set()
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                out["sm_max_mhz"] = float(parts[1])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            sm.sort()
            out["sm_mhz"] = sm[len(sm) // 2]
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def workload_cfg(name, rank, world):
    from llm_slo_ebpf_toolkit_b200 import synth
    if name == "c2":
        return synth.c2(rank=rank, world=world), "C2: 10M kernel events x 1M OTel spans, 64 pods / 8 nodes, 9 signal types, single-fault incidents (BASELINE.json configs[1])"
    if name == "c3":
        return synth.c3(rank=rank, world=world), "C3: 1B events x 50M spans, 4096 pods, multi-fault incidents (BASELINE.json configs[2])"
    if name == "c3s":   # C3's shape at 1/10 of its size
        return synth.config(100_000_000, 5_000_000, 4096, 16, 2, 10, 1, horizon_ns=360 * SECOND, rank=rank, world=world), \
            "C3/10: 100M events x 5M spans, 4096 pods, multi-fault incidents"
    if name == "small":
        return synth.config(1_000_000, 100_000, 64, 8, 2, 10, 0, horizon_ns=360 * SECOND, rank=rank, world=world), \
            "C2/10: 1M events x 100k spans (smoke-sized)"
    raise SystemExit("unknown workload %r" % name)


def sample_cfg(n_events, n_spans, multi=0):
    """A bounded sample of C2 with C2's density (events per pod-second): the horizon shrinks with the sample."""
    from llm_slo_ebpf_toolkit_b200 import synth
    horizon = max(int(3600 * SECOND * n_spans / 1_000_000), 4 * SECOND)
    return synth.config(n_events, n_spans, 64, 8, 2, 10, multi, horizon_ns=horizon)


def cpu_pass(w, brute, nthreads):
    """The whole path on the CPU oracle (test infrastructure used here only as the timed baseline)."""
    from oracle import cpu as O
    j = O.process_batch(w["spans"], w["events"], w["idents"], indexed=not brute, nthreads=nthreads)
    ni = len(w["inc_actual"])
    sig = O.incident_features(j["attr"], w["spans"]["incident"], ni)
    post, top1, _ = O.attribute(O.default_model(), sig)
    return O.confusion(w["inc_actual"], top1, w["inc_expected"], w["inc_expected_n"], post, 0.10, 0.01)


def run_reference(args, out):
    """--impl reference: the reference's own algorithm for the path — every span against every signal
    (ebpfcorrelator/processor.go:29-49, correlator.go:72-90), then AttributeSample and BuildConfusionMatrix —
    as restated in oracle/oracle.cpp (Go is not installed, so the reference itself cannot run), on all host
    threads, on a bounded sample of C2."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from llm_slo_ebpf_toolkit_b200 import synth
    from oracle import cpu as O
    O.build()
    threads = max(1, O.hw_threads())
    n_events, n_spans = 200_000, 20_000
    cfg = sample_cfg(n_events, n_spans)
    w = synth.generate_host(cfg)
    t_probe = time.perf_counter()
    cpu_pass(w, True, threads)                      # also the first warm-up step
    t_probe = time.perf_counter() - t_probe
    # keep the whole run within a few minutes whatever --steps is: the loop costs O(S x E), so shrink both sides
    budget_s = 150.0
    total_steps = max(args.steps + max(args.warmup - 1, 0), 1)
    for _ in range(3):   # the cost is not exactly quadratic (fixed per-pass overheads): re-probe after shrinking
        if t_probe * total_steps <= budget_s or n_spans <= 1_000:
            break
        scale = max((budget_s / (t_probe * total_steps)) ** 0.5 * 0.9, 0.05)
        n_events, n_spans = max(int(n_events * scale), 10_000), max(int(n_spans * scale), 1_000)
        cfg = sample_cfg(n_events, n_spans)
        w = synth.generate_host(cfg)
        t_probe = time.perf_counter()
        cpu_pass(w, True, threads)
        t_probe = time.perf_counter() - t_probe
    for _ in range(max(args.warmup - 1, 0)):
        cpu_pass(w, True, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        conf = cpu_pass(w, True, threads)
    dt = time.perf_counter() - t0
    value = n_events * args.steps / dt
    sample = ("brute force O(S x E) as the reference's ProcessBatch, on %d events x %d spans with C2's per-pod density "
              "(1/%d of C2); the loop's cost per event grows with S, so at full C2 (1M spans) the same code is ~%dx "
              "slower per event" % (n_events, n_spans, round(1_000_000 / n_spans), round(1_000_000 / n_spans)))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int64 keys / f64 scores", "data": "synthetic",
        "config": {"workload": workload_cfg("c2", 0, 1)[1], "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "checks": {"incidents": int(conf.n), "correct": int(conf.correct)},
    }
    out.write(json.dumps(line) + "\n")
    out.flush()
    return 0


def _claim_stdout():
    """Libraries (NCCL prints its version) write to fd 1; the contract is ONE JSON line there.  Keep a private handle
    on the real stdout for that line and point fd 1 at stderr for everything else."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


def main():
    out = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-buffer measurement (default min(steps, 10))")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-key-hint", action="store_true",
                    help="do not state the spans' time range (lsg_pipeline_args.key_hint): the step then reads the key layout back")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer measurement (C3 needs 37 GB of pinned memory)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args, out)
    args.warmup = max(args.warmup, 3)

    import numpy as np
    import torch
    import torch.distributed as dist

    import llm_slo_ebpf_toolkit_b200 as pkg
    from llm_slo_ebpf_toolkit_b200 import _abi as A
    from llm_slo_ebpf_toolkit_b200 import sharding
    from llm_slo_ebpf_toolkit_b200.device import DeviceWorkload

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: this path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.Stream()
    ctx = pkg.Context(local, stream.cuda_stream)
    cfg, wl_name = workload_cfg(args.workload, rank, world)

    def barrier():
        stream.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    with torch.cuda.stream(stream):
        wl = DeviceWorkload(ctx, cfg)
        if not args.no_key_hint:
            wl.state_key_range()   # a replay knows its window: no readback of the key layout inside the step

        # The only exchange of the path is the sum of the counter block (N > 1).  Two blocks alternate, so the allreduce of
        # batch k (asynchronous, on NCCL's stream) overlaps the kernels of batch k + 1; it is waited for before its
        # block is reused, and every batch's reduction completes inside the timed region.
        blocks = [wl.counters, torch.zeros_like(wl.counters)]
        pending = [None, None]
        state = {"i": 0}

        def step():
            b = state["i"] & 1
            state["i"] += 1
            if pending[b] is not None:
                pending[b].wait()
                pending[b] = None
            wl.args.counters_dev = blocks[b].data_ptr()
            wl.step()
            if world > 1:
                pending[b] = dist.all_reduce(blocks[b], op=dist.ReduceOp.SUM, async_op=True)

        def drain():
            for b in (0, 1):
                if pending[b] is not None:
                    pending[b].wait()
                    pending[b] = None
            wl.counters = blocks[(state["i"] - 1) & 1]   # the most recent batch's (reduced) block

        for _ in range(args.warmup):
            step()
        drain()
        barrier()
        sort_info = ctx.last_sort_info()
        launches0 = ctx.launch_count()
        vis = [x for x in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if x.strip()]
        sampler = ClockSampler(vis[local] if local < len(vis) else local)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        for _ in range(args.steps):
            step()
        drain()
        ev1.record(stream)
        barrier()
        clocks = sampler.stop()
        ms = ev0.elapsed_time(ev1)
        launches = ctx.launch_count() - launches0
        # per-kernel durations: the same steps again with a CUDA-event pair around every kernel.  The pairs cost 5 % of
        # the step (26 event records), so they are kept out of the region `value` is taken from.
        prof_steps = min(args.steps, 50)
        ctx.profile_read(reset=True)
        ctx.profile_enable(True)
        evp0, evp1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        evp0.record(stream)
        for _ in range(prof_steps):
            step()
        drain()
        evp1.record(stream)
        barrier()
        prof = ctx.profile_read(reset=True)
        ctx.profile_enable(False)
        prof_ms = evp0.elapsed_time(evp1)
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_max = float(t.item())
        counters = wl.read_counters()


    def measure_e2e():
        """Host (pinned) buffers through the public host API: every step uploads that step's records and downloads
        every per-span / per-incident result inside the timed region.  `value` uses the double-buffered form
        (lsg_pipeline_submit / lsg_pipeline_wait, two batches in flight, so upload, kernels and download overlap) with
        the per-span attributes in their compact form (kept candidates' event, tier, signal, value: the dense
        12-column row is max-per-signal of those); the dense form and one blocking lsg_pipeline call are timed beside it."""
        with torch.cuda.stream(stream):
            e2e_steps = args.e2e_steps or min(args.steps, 20)
            f = 3

            def pinned(nbytes):
                return torch.empty(max(int(nbytes), 1), dtype=torch.uint8, pin_memory=True)

            slots = []
            for _slot in range(2):
                host = {}
                for name in ("events", "spans", "idents", "inc_actual", "inc_expected", "inc_expected_n"):
                    src = getattr(wl, name)
                    buf = pinned(src.numel() * src.element_size())
                    buf.copy_(src.view(torch.uint8).cpu())
                    host[name] = buf
                outs = {"cand_event": pinned(wl.n_spans * f * 4), "cand_tier": pinned(wl.n_spans * f),
                        "attr": pinned(wl.n_spans * A.NSIG * 8), "conf": pinned(wl.n_spans * 8),
                        "cand_value": pinned(wl.n_spans * f * 8), "cand_sig": pinned(wl.n_spans * f),
                        "inc_signals": pinned(wl.n_incidents * A.NSIG * 8), "posterior": pinned(wl.n_incidents * A.NDOM * 8),
                        "top1": pinned(wl.n_incidents), "counters": pinned(C.sizeof(A.Counters))}

                def make(dense):
                    return A.PipelineArgs(host["spans"].data_ptr(), wl.n_spans, host["events"].data_ptr(), wl.n_events,
                                          host["idents"].data_ptr(), wl.n_idents, wl.n_incidents, host["inc_actual"].data_ptr(),
                                          host["inc_expected"].data_ptr(), host["inc_expected_n"].data_ptr(), 0, 0.0, f, 0.10,
                                          outs["cand_event"].data_ptr(), outs["cand_tier"].data_ptr(),
                                          outs["attr"].data_ptr() if dense else None, outs["conf"].data_ptr() if dense else None,
                                          outs["inc_signals"].data_ptr(), outs["posterior"].data_ptr(), outs["top1"].data_ptr(),
                                          outs["counters"].data_ptr(), None if dense else outs["cand_value"].data_ptr(),
                                          None if dense else outs["cand_sig"].data_ptr())
                forms = {"dense": make(True), "compact": make(False)}
                for a_ in forms.values():   # the same statement about the batch as in the device-resident steps
                    a_.key_hint, a_.max_grp_hint = wl.args.key_hint, wl.args.max_grp_hint
                    a_.span_ts_lo, a_.span_ts_hi = wl.args.span_ts_lo, wl.args.span_ts_hi
                slots.append((host, outs, forms))
            h2d = sum(slots[0][0][k].numel() for k in slots[0][0])
            common = ("cand_event", "cand_tier", "inc_signals", "posterior", "top1", "counters")
            d2h = {"dense": sum(slots[0][1][k].numel() for k in common + ("attr", "conf")),
                   "compact": sum(slots[0][1][k].numel() for k in common + ("cand_value", "cand_sig"))}
            hc = A.Counters()
            ok = True

            def blocking(form):
                nonlocal ok
                for _ in range(2):
                    ctx.call_dev("lsg_pipeline", C.byref(slots[0][2][form]), C.byref(hc))
                barrier()
                t0 = time.perf_counter()
                for _ in range(e2e_steps):
                    ctx.call_dev("lsg_pipeline", C.byref(slots[0][2][form]), C.byref(hc))
                barrier()
                ok = ok and (bytes(hc.confusion) == bytes(counters.confusion) if world == 1 else True)
                return (time.perf_counter() - t0) * 1e3

            def pipelined(form):
                nonlocal ok
                for s_ in range(2):
                    ctx.call_dev("lsg_pipeline_submit", C.byref(slots[s_][2][form]), s_)
                for s_ in range(2):
                    ctx.call_dev("lsg_pipeline_wait", s_, C.byref(hc))
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                t_wall = time.perf_counter()
                for step_i in range(e2e_steps):
                    s_ = step_i & 1
                    if step_i >= 2:
                        ctx.call_dev("lsg_pipeline_wait", s_, C.byref(hc))
                    ctx.call_dev("lsg_pipeline_submit", C.byref(slots[s_][2][form]), s_)
                    sharding.allreduce_counters(wl.counters)
                for step_i in range(max(e2e_steps - 2, 0), e2e_steps):
                    ctx.call_dev("lsg_pipeline_wait", step_i & 1, C.byref(hc))
                e1.record(stream)
                barrier()
                t_wall = (time.perf_counter() - t_wall) * 1e3
                ok = ok and (bytes(hc.confusion) == bytes(counters.confusion) if world == 1 else True)
                # the downloads finish on their own stream after the last kernel: the wall clock around the waits is the honest figure
                ms_ = max(e0.elapsed_time(e1), t_wall)
                t = torch.tensor([ms_], dtype=torch.float64, device="cuda")
                if world > 1:
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                return float(t.item())

            sync_ms = blocking("dense")
            dense_ms = pipelined("dense")
            e2e_ms = pipelined("compact")
        rate = lambda ms_: wl.n_events * world * e2e_steps / (ms_ / 1e3)
        return {"value": rate(e2e_ms), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h["compact"], "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps,
                "api": "lsg_pipeline_submit/lsg_pipeline_wait, 2 batches in flight, pinned host buffers; per-span attributes "
                       "as (signal, value) of the kept candidates (cand_sig / cand_value), dense rows not requested",
                "dense_outputs": {"value": rate(dense_ms), "ms_per_step": dense_ms / e2e_steps,
                                  "d2h_bytes_per_step": d2h["dense"],
                                  "note": "same calls with the dense 12-column attribute rows and the confidence column downloaded"},
                "sync_ms_per_step": sync_ms / e2e_steps, "sync_value": rate(sync_ms),
                "matches_device_run": bool(ok)}

    e2e = None if args.no_e2e else measure_e2e()

    f = 3
    n_events_total = wl.n_events * world
    value = n_events_total * args.steps / (ms_max / 1e3)

    # ---- roofline of the dominant kernel -------------------------------------------------------------
    peak, peak_src = measured_peak_gbs()
    key_bytes = 4 if sort_info["key_bits"] <= 31 else 8
    n, s_, i_ = wl.n_events, wl.n_spans, wl.n_incidents
    counting = "event_keys" not in prof or prof["event_keys"][1] == 0   # counting sort (small key space) or radix passes
    alg = {   # algorithmic bytes per launch: every operand read once, every result written once
        "event_stats": 32 * n, "event_keys": (32 + key_bytes) * n,
        "radix_hist": (32 + 8) * n if counting else key_bytes * n,                  # count keys: the records once, (key, rank) out
        "radix_scatter": (8 + 4) * n if counting else 2 * (key_bytes + 4) * n,      # place: (key, rank) in, index out
        "gather_events": (4 + 32 + 32) * n,
        "join": 32 * n + 4 * n + 32 * s_ + 32 * s_ + s_ * (f * 5 + A.NSIG * 8 + 8),   # records + index once, span, ranges, outputs
        "span_ranges": 32 * s_ + 32 * s_,
        "incident_features": A.NSIG * 8 * s_ + 4 * s_ + 2 * A.NSIG * 8 * i_, "attribute": (A.NSIG * 8 + A.NDOM * 8 + 1) * i_,
        "confusion": (A.NDOM * 8 + 7) * i_,
    }
    per_kernel = {}
    for name, (tot_ms, cnt) in prof.items():
        if cnt:
            per_kernel[name] = {"ms_per_step": tot_ms / prof_steps, "launches_per_step": cnt / prof_steps,
                                "ms_per_launch": tot_ms / cnt}
            if name in alg:
                per_kernel[name]["GBps"] = alg[name] / (tot_ms / cnt * 1e-3) / 1e9
    dom = max(per_kernel, key=lambda k: per_kernel[k]["ms_per_step"]) if per_kernel else None
    roofline = None
    if dom and dom in alg:
        achieved = per_kernel[dom]["GBps"]
        # DRAM bytes per launch of that kernel from the committed ncu --set full capture of this workload, if there is one
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "r01_ncu_traffic_%s.json" % args.workload)) as fh:
                cap = json.load(fh)["kernels"]
            sass = {"join": "k_join_tiles", "radix_scatter": "k_radix_scatter", "radix_hist": "k_radix_hist",
                    "gather_events": "k_gather_events_soa"}.get(dom)
            if sass in cap:
                traffic = cap[sass]["dram_bytes_per_launch"]
        except Exception:
            traffic = None
        roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": alg[dom], "ms_per_launch": per_kernel[dom]["ms_per_launch"]}
    pipeline_alg = wl.algorithmic_bytes()
    end_to_end_frac = (pipeline_alg * args.steps / (ms_max / 1e3) / 1e9) / peak

    # ---- CPU baseline on the host cores (rank 0, N = 1) ---------------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from llm_slo_ebpf_toolkit_b200 import synth
        from oracle import cpu as O
        threads = max(1, O.hw_threads())
        ne, ns = 2_000_000, 200_000
        w = synth.generate_host(sample_cfg(ne, ns))
        cpu_pass(w, False, threads)
        reps, t0 = 0, time.perf_counter()
        while reps < 3 or time.perf_counter() - t0 < 10.0:
            cpu_pass(w, False, threads)
            reps += 1
            if time.perf_counter() - t0 > 25.0:
                break
        dt = time.perf_counter() - t0
        cpu_baseline = {"value": ne * reps / dt, "unit": UNIT, "cores": threads, "kind": "port",
                        "sample": "indexed CPU oracle (sort + windowed search, proven equal to the reference's brute force) "
                                  "on %d events x %d spans at C2's density, %d passes; the literal brute force is timed by "
                                  "--impl reference" % (ne, ns, reps)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int64 keys / f64 scores", "data": "synthetic",
            "config": {"workload": wl_name, "events_per_gpu": wl.n_events, "spans_per_gpu": wl.n_spans,
                       "incidents_per_gpu": wl.n_incidents, "window_ms": 2000, "threshold": 0.7, "fanout": 3,
                       "l2": "inputs larger than L2 (%.0f MB of records per step)" % ((wl.n_events + wl.n_spans) * 32 / 1e6),
                       "sort": dict(sort_info, method="counting" if counting else "radix"),
                       "key_hint": "spans' time range and group count stated by the caller (checked on the device; --no-key-hint reads the key layout back instead)" if not args.no_key_hint else "none",
                       "parallelism": "shard by pod identity x%d, sum-allreduce of %d counters" % (world, A.COUNTER_WORDS)},
            "e2e": e2e,
            "gpu_launches": int(launches),
            "roofline": roofline,
            "pipeline_roofline": {"algorithmic_bytes_per_event": pipeline_alg / wl.n_events, "frac_of_peak": end_to_end_frac,
                                  "note": "whole step vs SURVEY.md §8(d)'s 96+152r+163q B/event"},
            "kernels": per_kernel,
            "kernel_timing": {"steps": prof_steps, "ms_per_step": prof_ms / prof_steps,
                              "note": "second pass over the same steps with a CUDA-event pair around every kernel; "
                                      "`value` comes from the first pass, without the pairs"},
            "cpu_baseline": cpu_baseline,
            "clocks": clocks,
            "checks": {"incidents": int(counters.confusion.n), "correct": int(counters.confusion.correct),
                       "candidates": int(counters.candidates), "kept": int(counters.kept),
                       "low_confidence": int(counters.debug.low_confidence)},
        }
        out.write(json.dumps(line) + "\n")
        out.flush()
    if world > 1:
        dist.destroy_process_group()
    ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
