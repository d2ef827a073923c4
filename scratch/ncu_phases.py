"""Aggregate an ncu source-page export (cuda,sass correlated) by source line / by the `// ---- ` phase markers of one file.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv; python scratch/ncu_phases.py src.csv [top] [file]
(file: join_slab.cuh by default, join.cu for the default join)"""
import collections
import csv
import re
import sys

path = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
rows = list(csv.reader(open(path)))
main_file = sys.argv[3] if len(sys.argv) > 3 else "join_slab.cuh"
src = open("llm-slo-ebpf-toolkit_b200/csrc/" + main_file).read().splitlines()
marks = [(i + 1, ln.strip()[8:60]) for i, ln in enumerate(src) if re.match(r"\s*// ---- ", ln)]
cur, hdr = None, None
line_agg = collections.defaultdict(lambda: [0, 0, 0, ""])
ph_agg = collections.defaultdict(lambda: [0, 0, 0])
ph_st = collections.defaultdict(collections.Counter)
ST = ("stall_barrier", "stall_long_sb", "stall_short_sb", "stall_wait", "stall_branch_resolving", "stall_math",
      "stall_not_selected", "stall_selected", "stall_mio", "stall_lg", "stall_no_inst")
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        ii, it, isamp = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
        ist = {n: hdr.index(n) for n in ST}
        continue
    if hdr is None:
        continue
    try:
        line = int(r[0])
        inst, thr, smp = int(r[ii]), int(r[it]), int(r[isamp])
    except Exception:
        continue
    a = line_agg[(cur, line)]
    a[0] += inst; a[1] += thr; a[2] += smp; a[3] = r[1][:80]
    if cur == main_file:
        ph = "start"
        for ln, name in marks:
            if ln <= line:
                ph = "%d %s" % (ln, name)
    else:
        ph = "inlined:" + cur
    b = ph_agg[ph]
    b[0] += inst; b[1] += thr; b[2] += smp
    for n, ix in ist.items():
        try:
            ph_st[ph][n] += int(r[ix])
        except Exception:
            pass
tot = sum(a[0] for a in line_agg.values()) or 1
ts = sum(a[2] for a in line_agg.values()) or 1
print("total warp instructions %d, samples %d" % (tot, ts))
for ph, a in sorted(ph_agg.items(), key=lambda kv: -kv[1][2]):
    print("%-62s inst %5.1f%% lanes %4.1f samples %5.1f%%  %s" % (ph[:62], 100 * a[0] / tot, a[1] / max(a[0], 1), 100 * a[2] / ts,
                                                                 dict(ph_st[ph].most_common(3))))
print()
for (f, l), a in sorted(line_agg.items(), key=lambda kv: -kv[1][2])[:top_n]:
    print("%-16s %4d inst %5.1f%% lanes %4.1f samples %5.1f%%  %s" % (f, l, 100 * a[0] / tot, a[1] / max(a[0], 1), 100 * a[2] / ts, a[3]))
