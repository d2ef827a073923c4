"""bench.py without a GPU: the pieces the driver depends on that need no device.

- `--impl reference` (the reference arm): one JSON line with the contract's keys, timed on the CPU
  oracle only (the product library is not mapped into that process), zero GPU launches.
- the algorithmic-byte table behind `roofline.achieved`: the join's 475 MB at C2.
- every workload name resolves, for every rank of a world, to a shard of the stated size.
- without a CUDA device the default arm fails loudly instead of falling back to the CPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout=300):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, cwd=ROOT, env=env, capture_output=True,
                          text=True, timeout=timeout)


def test_reference_arm_prints_one_contract_line_from_the_oracle_only():
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "1"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "correlated_and_attributed_events_per_sec" and d["unit"] == "events/s"
    assert d["higher_is_better"] is True and d["steps"] == 1 and d["n_gpus"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert d["gpu_launches"] == 0
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    # the brute force is O(S x E): the line says what the figure means at the full configuration
    assert d["pair_checks_per_s"] > d["value"] and 0 < d["events_per_s_at_full_c2"] < d["value"]
    assert d["checks"]["incidents"] > 0 and 0 < d["checks"]["correct"] <= d["checks"]["incidents"]


def test_default_arm_without_a_gpu_fails_loudly():
    r = _run(["--steps", "1", "--no-extra", "--no-e2e", "--no-cpu-baseline"], timeout=120)
    assert r.returncode != 0
    assert "no CPU fallback" in (r.stderr + r.stdout)
    assert not [ln for ln in r.stdout.splitlines() if ln.strip().startswith("{")]


def test_algorithmic_bytes_and_workloads():
    sys.path.insert(0, ROOT)
    import bench
    from llm_slo_ebpf_toolkit_b200 import _abi as A
    alg = bench.algorithmic_bytes(A, 10_000_000, 1_000_000, 100_000, 3, 4, two_level=False)
    assert alg["join"] == 475_000_000                       # 32 B/event + 36 B/span + 119 B/span of outputs
    assert alg["radix_hist"] == 400_000_000 and alg["radix_scatter"] == 120_000_000   # count (record + key, rank) and place
    assert all(v > 0 for v in alg.values())
    two = bench.algorithmic_bytes(A, 10_000_000, 1_000_000, 100_000, 3, 4, two_level=True)
    assert two["bucket_scatter"] > alg["bucket_scatter"]
    sizes = {"c2": (10_000_000, 1_000_000), "c3": (1_000_000_000, 50_000_000), "c3s": (100_000_000, 5_000_000),
             "c2g1": (10_000_000, 1_000_000), "c2fifth": (2_000_000, 200_000), "small": (1_000_000, 100_000)}
    for name, (n_ev, n_sp) in sizes.items():
        for rank, world in ((0, 1), (3, 8)):
            cfg, label = bench.workload_cfg(name, rank, world)
            assert cfg.n_events == n_ev and cfg.n_spans == n_sp and label
