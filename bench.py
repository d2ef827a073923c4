#!/usr/bin/env python3
"""bench.py — correlated + attributed kernel events per second on N x B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's algorithm on the host cores

A step = one pass of the whole hot path over one batch of synthetic input:
partition events and spans into (group, time slab) buckets -> slab join (Match, top-3, per-attribute max, counters)
-> per-incident features -> Bayesian attribution -> confusion matrix [-> sum-allreduce of the counter block, N > 1].
Workload of the headline line at N = 1: BASELINE.json configs[1] ("C2": 10 M kernel events x 1 M spans, 64 pods, 9
signal types, single-fault incidents); at N > 1 every rank holds its own C2-sized shard of distinct pods (weak scaling,
the partition key is the pod/node identity so the join is shard-local).

`value`      events/s with the inputs resident in HBM (device events around K steps, max over ranks)
`e2e`        the same metric through lsg_pipeline_submit/_wait with HOST (pinned) buffers: H2D of the records and D2H
             of every per-span / per-incident result inside the timed region
`roofline`   the dominant kernel's algorithmic bytes per launch / its mean duration (CUDA events recorded on the
             launch stream) against MEASURED_PEAKS.json
`cpu_baseline` the CPU oracle (indexed variant, all host threads) on a bounded sample of the same workload
`configs`    the other configurations BASELINE.json names, measured in the same process: C3 (1 B events x 50 M spans
             on one GPU; at N > 1 one C3-sized shard per rank = C4's shape) and the attribution-only leg (config 5's
             size: 40 M incidents through lsg_attribute_dev + lsg_confusion_dev)
`checks`     computed on the device results of the timed batch: counter invariants, three sampled join groups re-run
             on the indexed CPU oracle, and at N > 1 the allreduced counter block against the sum of the ranks' own
"""
import argparse
import ctypes as C
import json
import os
import shutil
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
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe).  Started before the
    warm-up so that the first samples exist when a 10 ms timed region begins; only samples inside [mark, stop] count."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.t_mark = None
        self.n_mark = 0
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def _lines(self):
        self.f.flush()
        with open(self.f.name) as fh:
            return fh.read().splitlines()

    def mark(self):
        """The timed region starts now: samples taken before this are the warm-up's."""
        if self.p is not None:
            self.n_mark = len(self._lines())

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.05)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        lines = self._lines()
        timed = lines[self.n_mark:]
        if len(timed) < 3:      # a region shorter than a few sampling periods: the samples just before it count too
            timed = lines[max(0, self.n_mark - 5):]
            out["note"] = "timed region shorter than the sampling period: includes the last warm-up samples"
        sm, reasons = [], set()
        for line in timed:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[1]))
                out["sm_max_mhz"] = float(parts[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[4:8]):
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


_T0 = time.perf_counter()


def progress(env, msg):
    """One line on stderr per stage (rank 0, or every rank with LSG_BENCH_TRACE=1): where a run that never prints its JSON
    line stopped."""
    if env.get("rank", 0) == 0 or os.environ.get("LSG_BENCH_TRACE"):
        sys.stderr.write("[bench %7.1fs rank %d] %s\n" % (time.perf_counter() - _T0, env.get("rank", 0), msg))
        sys.stderr.flush()


def workload_cfg(name, rank, world):
    from llm_slo_ebpf_toolkit_b200 import synth
    if name == "c2":
        return synth.c2(rank=rank, world=world), "C2: 10M kernel events x 1M OTel spans, 64 pods / 8 nodes, 9 signal types, single-fault incidents (BASELINE.json configs[1])"
    if name == "c3":
        return synth.c3(rank=rank, world=world), "C3: 1B events x 50M spans, 4096 pods, multi-fault incidents (BASELINE.json configs[2])"
    if name == "c3s":   # C3's shape at 1/10 of its size
        return synth.config(100_000_000, 5_000_000, 4096, 16, 2, 10, 1, horizon_ns=360 * SECOND, rank=rank, world=world), \
            "C3/10: 100M events x 5M spans, 4096 pods, multi-fault incidents"
    if name == "c2g1":   # C2's size with every pod in ONE join group: what a trace id that spans all pods does to the grouping
        return synth.config(10_000_000, 1_000_000, 64, 64, 64, 10, 0, rank=rank, world=world), \
            "C2 size, one join group: 10M events x 1M spans, 64 pods in a single group (every span's window exceeds a tile)"
    if name == "c2fifth":   # C2's density at 1/5 of its size: the 64 MB of event records stay in L2 (what the join costs without DRAM)
        return synth.config(2_000_000, 200_000, 64, 8, 2, 10, 0, horizon_ns=720 * SECOND, rank=rank, world=world), \
            "C2/5: 2M events x 200k spans at C2's density (event records L2-resident)"
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
    as restated in oracle/oracle.cpp, on all host threads, on a bounded sample of C2.  The reference is a Go module: no
    Go toolchain exists in the build image and /root/reference does not travel to the GPU box, so the unmodified
    reference cannot run there even where `go` is installed (the probe's answer is printed as `go`).  This arm maps
    oracle/liboracle.so only: its inputs come from the host-only copy of the generator inside that library."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import cpu as O
    O.build()
    threads = max(1, O.hw_threads())
    n_events, n_spans = 200_000, 20_000
    w = O.generate_host(sample_cfg(n_events, n_spans))
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
        w = O.generate_host(sample_cfg(n_events, n_spans))
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
    pair_checks = float(n_events) * n_spans * args.steps / dt
    sample = ("brute force O(S x E) as the reference's ProcessBatch, on %d events x %d spans with C2's per-pod density "
              "(1/%d of C2); the loop's cost per event grows with S, so at full C2 (1M spans) the same code is ~%dx "
              "slower per event" % (n_events, n_spans, round(1_000_000 / n_spans), round(1_000_000 / n_spans)))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int64 keys / f64 scores", "data": "synthetic",
        "config": {"workload": workload_cfg("c2", 0, 1)[1], "sample": sample, "same_config": False},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        # what `value` means on another size: the loop does pair_checks_per_s Match() calls per second whatever the
        # batch is, so at full C2 (1 M spans per event) it handles pair_checks_per_s / 1e6 events per second
        "pair_checks_per_s": pair_checks, "events_per_s_at_full_c2": pair_checks / 1_000_000,
        "go": shutil.which("go"), "reference_sources_on_this_box": os.path.isdir("/root/reference"),
        "gpu_launches": 0, "checks": {"incidents": int(conf.n), "correct": int(conf.correct)},
    }
    out.write(json.dumps(line) + "\n")
    out.flush()
    return 0


def numa_bind(cuda_index):
    """Run this rank on the CPUs of its GPU's NUMA node, before any pinned buffer is allocated (first touch places the
    pages on that node): with eight ranks uploading at once, buffers on the far socket halve a GPU's PCIe rate.
    Returns what was done, for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = [x for x in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if x.strip()]
        h = pynvml.nvmlDeviceGetHandleByIndex(int(vis[cuda_index]) if cuda_index < len(vis) and vis[cuda_index].isdigit() else cuda_index)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dev = "/sys/bus/pci/devices/" + bus.lower()[-12:]
        with open(dev + "/numa_node") as f:
            node = int(f.read().strip())
        with open(dev + "/local_cpulist") as f:
            cpulist = f.read().strip()
        cpus = set()
        for part in cpulist.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        allowed = os.sched_getaffinity(0)
        use = (cpus & allowed) or allowed
        os.sched_setaffinity(0, use)
        return {"gpu_numa_node": node, "cpus": len(use), "pci": bus, "_all_cpus": sorted(allowed)}
    except Exception as e:   # no NVML / no sysfs: run where the launcher put us
        return {"unbound": "%s: %s" % (type(e).__name__, e)}


def _claim_stdout():
    """Libraries (NCCL prints its version) write to fd 1; the contract is ONE JSON line there.  Keep a private handle
    on the real stdout for that line and point fd 1 at stderr for everything else."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


def algorithmic_bytes(A, n, s_, i_, f, key_bytes, two_level):
    """Algorithmic bytes per launch group: every operand read once, every result written once."""
    lv = 2 if two_level else 1
    return {
        "event_stats": 32 * s_,
        "bucket_count": 32 * n,                               # the records once; the counters live in L2
        "bucket_scatter": lv * (32 + 32 + 4) * n + (lv - 1) * 4 * n,   # records in, records + input index out (per level)
        "bucket_spans": 32 * s_ + (32 + 32 + 4) * s_,          # count + scatter of the spans
        "join": 32 * n + (32 + 4) * s_ + s_ * (f * 5 + A.NSIG * 8 + 8),   # bucketed records once, spans, outputs
        "event_keys": (32 + key_bytes) * n, "radix_hist": (32 + 8) * n, "radix_scatter": (8 + 4) * n,
        "gather_events": (4 + 32 + 32) * n, "span_ranges": 32 * s_ + 32 * s_,
        "incident_features": A.NSIG * 8 * s_ + 4 * s_ + 2 * A.NSIG * 8 * i_, "attribute": (A.NSIG * 8 + A.NDOM * 8 + 1) * i_,
        "confusion": (A.NDOM * 8 + 7) * i_,
    }


def verify_batch(torch, wl, counters, A, n_groups=3):
    """Checks on the device results of the last batch (rank-local, before any reduction):
    invariants of the counter block, and `n_groups` whole join groups re-run on the indexed CPU oracle (every span of
    the group: kept candidates in order, tiers).  Returns (ok, details)."""
    import numpy as np
    from oracle import cpu as O
    d, c = counters.debug, counters.confusion
    f = wl.fanout if wl.fanout > 0 else 3
    det = {}
    matrix = np.array([[c.matrix[a][p] for p in range(A.NDOM)] for a in range(A.MAX_ACTUAL)], dtype=np.int64)
    det["confusion_sums_to_incidents"] = int(matrix.sum()) == int(c.n) == wl.n_incidents
    pairs = wl.n_events * wl.n_spans
    det["pairs_accounted"] = int(d.unmatched) + int(d.low_confidence) + int(counters.candidates) == pairs - int(d.unsupported_type)
    ce = wl.cand_event.view(torch.int32)[: wl.n_spans * f]
    kept_dev = int((ce != -1).sum().item())
    det["kept_equals_listed"] = kept_dev == int(counters.kept)
    det["fanout_dropped"] = int(d.fanout_dropped) == int(counters.candidates) - int(counters.kept)   # kept = min(candidates, fanout) per span
    # whole join groups on the oracle
    idents = wl.idents.cpu().numpy().view(A.IDENT_DTYPE)[: wl.n_idents]
    grps = np.unique(idents["grp"][idents["grp"] != 0])
    picks = sorted(set(int(grps[k]) for k in (0, len(grps) // 2, len(grps) - 1))) if len(grps) else []
    ev32 = wl.events.view(torch.int32).view(-1, 8)[: wl.n_events]
    sp32 = wl.spans.view(torch.int32).view(-1, 8)[: wl.n_spans]
    ct = wl.cand_tier[: wl.n_spans * f].view(-1, f)
    ce2 = ce.view(-1, f)
    groups_ok, n_sp_checked, n_ev_checked = True, 0, 0
    for g in picks[:n_groups]:
        rows = torch.from_numpy(np.nonzero(idents["grp"] == g)[0].astype(np.int32)).to(ev32.device)
        e_idx = torch.nonzero(torch.isin(ev32[:, 4], rows)).flatten()
        s_idx = torch.nonzero(torch.isin(sp32[:, 2], rows)).flatten()
        ev = ev32[e_idx].cpu().numpy().view(A.EVENT_DTYPE).reshape(-1)
        sp = sp32[s_idx].cpu().numpy().view(A.SPAN_DTYPE).reshape(-1)
        want = O.process_batch(sp, ev, idents, indexed=True, nthreads=max(1, O.hw_threads()))
        e_host = e_idx.cpu().numpy().astype(np.uint32)
        wce = want["cand_event"]
        wmap = np.where(wce == A.NONE, np.uint32(A.NONE), e_host[np.minimum(wce, max(len(e_host) - 1, 0))] if len(e_host) else np.uint32(A.NONE))
        got_e = ce2[s_idx].cpu().numpy().view(np.uint32)
        got_t = ct[s_idx].cpu().numpy()
        groups_ok = groups_ok and np.array_equal(got_e, wmap) and np.array_equal(got_t, want["cand_tier"])
        n_sp_checked += len(sp)
        n_ev_checked += len(ev)
    det["sampled_groups_equal_oracle"] = bool(groups_ok)
    det["sampled"] = {"groups": picks[:n_groups], "spans": n_sp_checked, "events": n_ev_checked}
    ok = all(v for k, v in det.items() if isinstance(v, bool))
    return ok, det


def device_leg(env, name, steps, warmup, key_hint=True, prof_steps_max=50, keep=False):
    """One workload with its inputs resident in HBM: warm-up, K timed steps (device events, max over ranks), a second
    pass with per-kernel event pairs, the checks.  Returns (result dict, workload or None)."""
    torch, dist, ctx, stream, A = env["torch"], env["dist"], env["ctx"], env["stream"], env["A"]
    rank, world = env["rank"], env["world"]
    from llm_slo_ebpf_toolkit_b200.device import DeviceWorkload
    cfg, wl_name = workload_cfg(name, rank, world)

    def barrier():
        stream.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    with torch.cuda.stream(stream):
        progress(env, "%s: generating the shard on the device" % name)
        wl = DeviceWorkload(ctx, cfg)
        if key_hint:
            wl.state_key_range()   # a replay knows its window: no readback of the key layout inside the step
        # The only exchange of the path is the sum of the counter block (N > 1).  Two blocks alternate, so the allreduce of
        # batch k (asynchronous, on NCCL's stream) overlaps the kernels of batch k + 1; it is waited for before its
        # block is reused, and every batch's reduction completes inside the timed region.
        blocks = [wl.counters, torch.zeros_like(wl.counters)]
        pending = [None, None]
        state = {"i": 0}

        def step(reduce=True):
            b = state["i"] & 1
            state["i"] += 1
            if pending[b] is not None:
                pending[b].wait()
                pending[b] = None
            wl.args.counters_dev = blocks[b].data_ptr()
            wl.step()
            if world > 1 and reduce:
                pending[b] = dist.all_reduce(blocks[b], op=dist.ReduceOp.SUM, async_op=True)
            return b

        def drain():
            for b in (0, 1):
                if pending[b] is not None:
                    pending[b].wait()
                    pending[b] = None

        progress(env, "%s: warm-up" % name)
        for _ in range(warmup):
            step()
        drain()
        barrier()
        progress(env, "%s: %d timed steps" % (name, steps))
        sort_info = ctx.last_sort_info()
        launches0 = ctx.launch_count()
        if env.get("sampler"):
            env["sampler"].mark()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        for _ in range(steps):
            last_block = step()
        drain()
        ev1.record(stream)
        barrier()
        ms = ev0.elapsed_time(ev1)
        launches = ctx.launch_count() - launches0
        reduced = blocks[last_block].clone()
        # per-kernel durations: the same steps again with a CUDA-event pair around every kernel.  The pairs cost a few
        # percent of the step, so they are kept out of the region `value` is taken from.
        prof_steps = min(steps, prof_steps_max)
        progress(env, "%s: %.3f ms/step on this rank; per-kernel pass" % (name, ms / steps))
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
        # ---- checks: one more batch without the reduction gives this rank's own counters
        progress(env, "%s: checks (counter invariants, sampled groups on the CPU oracle)" % name)
        b = step(reduce=False)
        barrier()
        wl.counters = blocks[b]
        local = wl.read_counters()
        ok, det = verify_batch(torch, wl, local, A)
        progress(env, "%s: checks done on this rank: %s" % (name, ok))
        if world > 1:
            # the allreduced block of the timed region must be the sum of the ranks' own blocks (same inputs every step)
            gathered = [torch.zeros_like(blocks[b]) for _ in range(world)]
            dist.all_gather(gathered, blocks[b])
            det["allreduce_equals_sum_of_ranks"] = bool(torch.equal(torch.stack(gathered).sum(0), reduced))
            ok = ok and det["allreduce_equals_sum_of_ranks"]
            flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            ok = bool(flag.item())
            wl.counters = reduced
            counters = wl.read_counters()
            wl.counters = blocks[b]
        else:
            counters = local

    f = 3
    n, s_, i_ = wl.n_events, wl.n_spans, wl.n_incidents
    peak, peak_src = measured_peak_gbs()
    key_bytes = 4 if sort_info["key_bits"] <= 31 else 8
    slab = "bucket_count" in prof and prof["bucket_count"][1] > 0
    n_buckets = ((1 << max(sort_info["key_bits"], 6)) >> 6)
    alg = algorithmic_bytes(A, n, s_, i_, f, key_bytes, two_level=slab and n_buckets > (1 << 17))
    per_kernel = {}
    for kname, (tot_ms, cnt) in prof.items():
        if cnt:
            per_kernel[kname] = {"ms_per_step": tot_ms / prof_steps, "launches_per_step": cnt / prof_steps,
                                 "ms_per_launch": tot_ms / cnt}
            if kname in alg:
                per_kernel[kname]["GBps"] = alg[kname] / (tot_ms / cnt * 1e-3) / 1e9
    # the spans' partition runs on a side stream next to the events': it is not on the step's critical path
    crit = {k: v for k, v in per_kernel.items() if k not in ("bucket_spans", "span_sort")}
    dom = max(crit, key=lambda k: crit[k]["ms_per_step"]) if crit else None
    roofline = None
    if dom and dom in alg:
        achieved = per_kernel[dom]["GBps"]
        traffic, cap_src = None, None
        try:   # DRAM bytes per launch of that kernel from the committed ncu --set full capture of this workload
            with open(os.path.join(ROOT, "profiles", "r02_ncu_traffic_%s.json" % name)) as fh:
                capj = json.load(fh)
            sass = ({"join": "k_join_slab", "bucket_scatter": "k_bucket_scatter", "bucket_count": "k_bucket_count"} if slab else
                    {"join": "k_join_tiles", "radix_hist": "k_count_keys" if "event_keys" not in per_kernel else "k_radix_hist",
                     "radix_scatter": "k_place_events" if "event_keys" not in per_kernel else "k_radix_scatter"}).get(dom)
            if sass == "k_join_tiles" and sass not in capj["kernels"] and "k_join_ring" in capj["kernels"]:
                sass = "k_join_ring"   # dense batches take the sliding-tile form of the kernel
            if sass in capj["kernels"]:
                traffic = capj["kernels"][sass]["dram_bytes_per_launch"]
                cap_src = "ncu --set full capture at commit %s (profiles/r02_ncu_traffic_%s.json)" % (capj.get("commit", "?"), name)
        except Exception:
            pass
        roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                    "frac": achieved / peak, "traffic": traffic, "traffic_source": cap_src, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": alg[dom], "ms_per_launch": per_kernel[dom]["ms_per_launch"]}
    pipeline_alg = wl.algorithmic_bytes()
    res = {
        "workload": wl_name, "value": n * world * steps / (ms_max / 1e3), "unit": UNIT,
        "incidents_per_s": i_ * world * steps / (ms_max / 1e3),
        "steps": steps, "warmup": warmup, "ms_per_step": ms_max / steps, "gpu_launches": int(launches),
        "events_per_gpu": n, "spans_per_gpu": s_, "incidents_per_gpu": i_,
        "sort": dict(sort_info, method="bucket partition + slab join" if slab else "permutation (counting/radix) + gathering join"),
        "roofline": roofline,
        "pipeline_roofline": {"algorithmic_bytes_per_event": pipeline_alg / n,
                              "frac_of_peak": (pipeline_alg * steps / (ms_max / 1e3) / 1e9) / peak,
                              "note": "whole step vs SURVEY.md §8(d)'s 96+152r+163q B/event"},
        "kernels": per_kernel,
        "kernel_timing": {"steps": prof_steps, "ms_per_step": prof_ms / prof_steps,
                          "note": "second pass over the same steps with a CUDA-event pair around every kernel; `value` "
                                  "comes from the first pass, without the pairs"},
        "checks": dict(det, verified=bool(ok), incidents=int(counters.confusion.n), correct=int(counters.confusion.correct),
                       candidates=int(counters.candidates), kept=int(counters.kept),
                       low_confidence=int(counters.debug.low_confidence), bad_idents=int(counters.bad_idents)),
    }
    env["local_counters"] = local
    if keep:
        return res, wl
    del wl, blocks
    torch.cuda.empty_cache()
    return res, None


def attribution_leg(env, n_inc, steps=5):
    """Attribution alone at config 5's size: lsg_attribute_dev + lsg_confusion_dev over n_inc incidents per GPU."""
    torch, ctx, stream, A, world, dist = env["torch"], env["ctx"], env["stream"], env["A"], env["world"], env["dist"]
    with torch.cuda.stream(stream):
        g = torch.Generator(device="cuda")
        g.manual_seed(42 + env["rank"])
        sig = torch.rand(n_inc, A.NSIG, dtype=torch.float64, device="cuda", generator=g) * 120.0
        absent = torch.rand(n_inc, A.NSIG, device="cuda", generator=g) < 0.25
        sig[absent] = float("nan")
        del absent
        post = torch.empty(n_inc, A.NDOM, dtype=torch.float64, device="cuda")
        top1 = torch.empty(n_inc, dtype=torch.uint8, device="cuda")
        actual = torch.randint(0, 8, (n_inc,), dtype=torch.uint8, device="cuda", generator=g)
        expected = actual.to(torch.int32)
        exp_n = torch.ones(n_inc, dtype=torch.uint8, device="cuda")
        out = A.ConfusionOut()

        def once():
            ctx.call_dev("lsg_attribute_dev", sig.data_ptr(), None, None, n_inc, post.data_ptr(), top1.data_ptr(), None)
            ctx.call_dev("lsg_confusion_dev", actual.data_ptr(), top1.data_ptr(), expected.data_ptr(), exp_n.data_ptr(),
                         post.data_ptr(), 0.10, 0.01, n_inc, C.byref(out))
        for _ in range(3):
            once()
        stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            once()
        e1.record(stream)
        stream.synchronize()
        ms = e0.elapsed_time(e1)
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        psum = post.sum(1)
        has = ~torch.isnan(sig).all(1)   # an incident with no signal at all has no posteriors (the rule path)
        psum_ok = bool(((psum[has] - 1.0).abs() < 1e-9).all().item()) and bool(torch.isnan(psum[~has]).all().item())
    peak, _ = measured_peak_gbs()
    alg = (A.NSIG * 8 + A.NDOM * 8 + 1) * n_inc + (A.NDOM * 8 + 7) * n_inc
    return {"workload": "attribution only: %d incidents per GPU (config 5's size over 8 GPUs = 40 M), 12 signals each, "
                        "25%% absent; lsg_attribute_dev + lsg_confusion_dev (each call synchronises)" % n_inc,
            "incidents_per_s": n_inc * world * steps / (ms / 1e3), "ms_per_step": ms / steps, "steps": steps,
            "hbm_frac": alg * steps / (ms / 1e3) / 1e9 / peak, "algorithmic_bytes_per_incident": alg / n_inc,
            "checks": {"posteriors_sum_to_one": psum_ok, "confusion_n": int(out.n) == n_inc,
                       "verified": bool(psum_ok and int(out.n) == n_inc)}}


def measure_e2e(env, wl, args):
    """Host (pinned) buffers through the public host API: every step uploads that step's records and downloads
    every per-span / per-incident result inside the timed region.  `value` uses the double-buffered form
    (lsg_pipeline_submit / lsg_pipeline_wait, two batches in flight, so upload, kernels and download overlap) with
    the per-span attributes in their compact form (kept candidates' event, tier, signal, value: the dense
    12-column row is max-per-signal of those); the dense form and one blocking lsg_pipeline call are timed beside it."""
    torch, dist, ctx, stream, A = env["torch"], env["dist"], env["ctx"], env["stream"], env["A"]
    world = env["world"]
    local = env["local_counters"]
    import llm_slo_ebpf_toolkit_b200 as pkg_mod
    pkg_lib = pkg_mod.load_library()

    def barrier():
        stream.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

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

            # the events once more in their 24-byte upload form (what an interner would emit; packed outside the timed region
            # like every other input of the step)
            import numpy as np
            ev_np = host["events"].numpy().view(A.EVENT_DTYPE)
            e24 = pinned(wl.n_events * 24)
            e24_base = int(wl.args.span_ts_lo) - 2_000_000_000 if wl.args.key_hint else int(wl.cfg.t0_ns) - 2_000_000_000
            misfit = pkg_lib.lsg_pack_events24(ev_np.ctypes.data, wl.n_events, e24_base, e24.numpy().ctypes.data)
            host["events24"] = e24

            def make(dense, compact_in=False):
                a = make32(dense)
                if compact_in and misfit == 0:
                    a.events, a.events24, a.events24_base = None, e24.data_ptr(), e24_base
                return a

            def make32(dense):
                return A.PipelineArgs(host["spans"].data_ptr(), wl.n_spans, host["events"].data_ptr(), wl.n_events,
                                      host["idents"].data_ptr(), wl.n_idents, wl.n_incidents, host["inc_actual"].data_ptr(),
                                      host["inc_expected"].data_ptr(), host["inc_expected_n"].data_ptr(), 0, 0.0, f, 0.10,
                                      outs["cand_event"].data_ptr(), outs["cand_tier"].data_ptr(),
                                      outs["attr"].data_ptr() if dense else None, outs["conf"].data_ptr() if dense else None,
                                      outs["inc_signals"].data_ptr(), outs["posterior"].data_ptr(), outs["top1"].data_ptr(),
                                      outs["counters"].data_ptr(), None if dense else outs["cand_value"].data_ptr(),
                                      None if dense else outs["cand_sig"].data_ptr())
            forms = {"dense": make(True), "compact": make(False), "compact24": make(False, True)}
            for a_ in forms.values():   # the same statement about the batch as in the device-resident steps
                a_.key_hint, a_.max_grp_hint = wl.args.key_hint, wl.args.max_grp_hint
                a_.span_ts_lo, a_.span_ts_hi = wl.args.span_ts_lo, wl.args.span_ts_hi
            slots.append((host, outs, forms))
        h2d = sum(slots[0][0][k].numel() for k in slots[0][0] if k != "events24")
        h2d24 = h2d - slots[0][0]["events"].numel() + slots[0][0]["events24"].numel()
        common = ("cand_event", "cand_tier", "inc_signals", "posterior", "top1", "counters")
        d2h = {"dense": sum(slots[0][1][k].numel() for k in common + ("attr", "conf")),
               "compact": sum(slots[0][1][k].numel() for k in common + ("cand_value", "cand_sig"))}
        hc = A.Counters()
        ok = True

        def same_as_device_run():
            # every rank compares what came back through the host API with its own device-resident run
            return bytes(hc.confusion) == bytes(local.confusion) and hc.debug.as_tuple() == local.debug.as_tuple() and \
                int(hc.kept) == int(local.kept)

        def blocking(form):
            nonlocal ok
            for _ in range(2):
                ctx.call_dev("lsg_pipeline", C.byref(slots[0][2][form]), C.byref(hc))
            barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                ctx.call_dev("lsg_pipeline", C.byref(slots[0][2][form]), C.byref(hc))
            barrier()
            ok = ok and same_as_device_run()
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
            blk = torch.zeros(A.COUNTER_WORDS, dtype=torch.int64, device="cuda")
            for step_i in range(e2e_steps):
                s_ = step_i & 1
                if step_i >= 2:
                    ctx.call_dev("lsg_pipeline_wait", s_, C.byref(hc))
                ctx.call_dev("lsg_pipeline_submit", C.byref(slots[s_][2][form]), s_)
                if world > 1:
                    dist.all_reduce(blk, op=dist.ReduceOp.SUM)     # the path's one exchange, once per batch
            for step_i in range(max(e2e_steps - 2, 0), e2e_steps):
                ctx.call_dev("lsg_pipeline_wait", step_i & 1, C.byref(hc))
            e1.record(stream)
            barrier()
            t_wall = (time.perf_counter() - t_wall) * 1e3
            ok = ok and same_as_device_run()
            # the downloads finish on their own stream after the last kernel: the wall clock around the waits is the honest figure
            ms_ = max(e0.elapsed_time(e1), t_wall)
            t = torch.tensor([ms_], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())

        sync_ms = blocking("dense")
        dense_ms = pipelined("dense")
        rec32_ms = pipelined("compact")
        e2e_ms = pipelined("compact24")
        if world > 1:
            flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            ok = bool(flag.item())
    rate = lambda ms_: wl.n_events * world * e2e_steps / (ms_ / 1e3)
    return {"value": rate(e2e_ms), "unit": UNIT, "h2d_bytes_per_step": h2d24 if misfit == 0 else h2d,
            "d2h_bytes_per_step": d2h["compact"], "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps,
            "incidents_per_s": wl.n_incidents * world * e2e_steps / (e2e_ms / 1e3),
            "api": "lsg_pipeline_submit/lsg_pipeline_wait, 2 batches in flight, pinned host buffers; events in the 24-byte "
                   "upload form (lsg_event24, widened by the first device pass), spans / identities / incidents as "
                   "they are; per-span attributes as (signal, value) of the kept candidates (cand_sig / cand_value), "
                   "dense rows not requested",
            "events24_misfit": int(misfit),
            "records32": {"value": rate(rec32_ms), "ms_per_step": rec32_ms / e2e_steps, "h2d_bytes_per_step": h2d,
                          "note": "same calls with the events uploaded as 32-byte lsg_event records"},
            "dense_outputs": {"value": rate(dense_ms), "ms_per_step": dense_ms / e2e_steps,
                              "d2h_bytes_per_step": d2h["dense"], "h2d_bytes_per_step": h2d,
                              "note": "32-byte events up; the dense 12-column attribute rows and the confidence column down"},
            "sync_ms_per_step": sync_ms / e2e_steps, "sync_value": rate(sync_ms),
            "matches_device_run": bool(ok)}


def main():
    out = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-buffer measurement (default min(steps, 20))")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-key-hint", action="store_true",
                    help="do not state the spans' time range (lsg_pipeline_args.key_hint): the step then reads the key layout back")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer measurement (C3 needs 37 GB of pinned memory)")
    ap.add_argument("--no-extra", action="store_true", help="skip the C3 / attribution-only legs (`configs`)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args, out)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist

    import llm_slo_ebpf_toolkit_b200 as pkg
    from llm_slo_ebpf_toolkit_b200 import _abi as A

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: this path has no CPU fallback")
    numa = numa_bind(local) if os.environ.get("LSG_NO_NUMA_BIND") is None else {"unbound": "LSG_NO_NUMA_BIND"}
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.Stream()
    ctx = pkg.Context(local, stream.cuda_stream)
    vis = [x for x in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if x.strip()]
    sampler = ClockSampler(vis[local] if local < len(vis) else local)   # started before the warm-up
    env = {"torch": torch, "dist": dist, "ctx": ctx, "stream": stream, "A": A, "rank": rank, "world": world,
           "sampler": sampler}

    main_res, wl = device_leg(env, args.workload, args.steps, args.warmup, key_hint=not args.no_key_hint, keep=True)
    clocks = sampler.stop()
    env["sampler"] = None
    progress(env, "device-resident leg done; end-to-end leg" if not args.no_e2e else "device-resident leg done")
    e2e = None if args.no_e2e else measure_e2e(env, wl, args)
    progress(env, "end-to-end leg done")
    n_events_main = wl.n_events
    del wl
    torch.cuda.empty_cache()

    # ---- CPU baseline on the host cores (rank 0, N = 1) ---------------------------------------------
    cpu_baseline = None
    all_cpus = numa.pop("_all_cpus", None)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import cpu as O
        if all_cpus:
            os.sched_setaffinity(0, all_cpus)   # the CPU baseline uses every host core, not only the GPU's node
        threads = max(1, O.hw_threads())
        ne, ns = 2_000_000, 200_000
        w = O.generate_host(sample_cfg(ne, ns))
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

    # ---- the one JSON line: built now, so that it survives an extra leg that hangs ---------------------------
    import threading
    r = main_res
    emit_lock = threading.Lock()
    emitted = {"done": False}

    def emit(configs):
        """Rank 0 prints the line, once."""
        with emit_lock:
            if emitted["done"]:
                return
            emitted["done"] = True
            if rank != 0:
                return
            line = {
                "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "int64 keys / f64 scores", "data": "synthetic",
                "incidents_per_s": r["incidents_per_s"],
                "config": {"workload": r["workload"], "events_per_gpu": r["events_per_gpu"], "spans_per_gpu": r["spans_per_gpu"],
                           "incidents_per_gpu": r["incidents_per_gpu"], "window_ms": 2000, "threshold": 0.7, "fanout": 3,
                           "l2": "inputs larger than L2 (%.0f MB of records per step)" % ((r["events_per_gpu"] + r["spans_per_gpu"]) * 32 / 1e6),
                           "sort": r["sort"],
                           "key_hint": "spans' time range and group count stated by the caller (checked on the device; --no-key-hint reads the key layout back instead)" if not args.no_key_hint else "none",
                           "parallelism": "shard by pod identity x%d, sum-allreduce of %d counters" % (world, A.COUNTER_WORDS),
                           "numa": numa},
                "e2e": e2e,
                "gpu_launches": r["gpu_launches"],
                "roofline": r["roofline"],
                "pipeline_roofline": r["pipeline_roofline"],
                "kernels": r["kernels"],
                "kernel_timing": r["kernel_timing"],
                "cpu_baseline": cpu_baseline,
                "clocks": clocks,
                "checks": r["checks"],
                "configs": configs,
            }
            out.write(json.dumps(line) + "\n")
            out.flush()

    # ---- the other configurations BASELINE.json names, in the same process -------------------------------
    configs = {}
    if not args.no_extra and args.workload == "c2":
        budget = float(os.environ.get("LSG_BENCH_EXTRA_BUDGET_S", "600"))

        def give_up():
            progress(env, "the extra legs did not finish within %.0f s: printing the line without them" % budget)
            emit(dict(configs, error="extra legs (C3 / attribution only) exceeded %.0f s and were abandoned" % budget))
            os._exit(0)

        watchdog = threading.Timer(budget, give_up)
        watchdog.daemon = True
        watchdog.start()
        try:
            free = torch.cuda.mem_get_info()[0]
            if free > 150 * (1 << 30):
                res, _ = device_leg(env, "c3", 5, 3, key_hint=not args.no_key_hint, prof_steps_max=3)
                configs["c3" if world == 1 else "c4_shape"] = dict(
                    res, note="BASELINE.json configs[2] on one GPU" if world == 1 else
                    "one C3-sized shard (1 B events x 50 M spans) per rank: configs[3]'s shape at %d GPUs" % world)
            else:
                configs["c3"] = {"skipped": "only %.0f GB of device memory free" % (free / 2**30)}
        except Exception as e:   # the headline line must survive a failure of an extra leg
            configs["c3"] = {"error": "%s: %s" % (type(e).__name__, e)}
        try:
            progress(env, "attribution-only leg")
            configs["attribution_only"] = attribution_leg(env, 40_000_000 // 8 if world > 1 else 40_000_000)
        except Exception as e:
            configs["attribution_only"] = {"error": "%s: %s" % (type(e).__name__, e)}
        watchdog.cancel()

    progress(env, "all legs done")
    emit(configs)
    if world > 1:
        dist.destroy_process_group()
    ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
