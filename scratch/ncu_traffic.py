"""ncu --page raw --csv of a `--set full` capture -> per-kernel DRAM bytes per launch and a few utilisation figures
(profiles/r01_ncu_traffic_<workload>.json, read by bench.py for roofline.traffic).
usage: ncu_traffic.py raw.csv workload "source note" [commit] > out.json"""
import csv, json, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
def f(r, name):
    try:
        return float(r[col[name]].replace(',', ''))
    except Exception:
        return None
def scale(name):
    u = units[col[name]].lower()
    return {'byte': 1.0, 'kbyte': 1e3, 'mbyte': 1e6, 'gbyte': 1e9}.get(u, 1.0)
def tscale(name):
    u = units[col[name]].lower()
    return {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6}.get(u, 1.0)
agg = {}
for r in rows[2:]:
    name = r[col['Kernel Name']].replace('void ', '').replace('<unnamed>::', '')
    name = re.split(r'[<(]', name)[0].strip()
    a = agg.setdefault(name, {'n': 0, 'rd': 0.0, 'wr': 0.0, 'us': 0.0, 'issue': 0.0, 'inst': 0.0, 'regs': 0, 'grid': 0})
    a['n'] += 1
    a['rd'] += f(r, 'dram__bytes_read.sum') * scale('dram__bytes_read.sum')
    a['wr'] += f(r, 'dram__bytes_write.sum') * scale('dram__bytes_write.sum')
    a['us'] += f(r, 'gpu__time_duration.sum') * tscale('gpu__time_duration.sum')
    a['issue'] += f(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active') or 0.0
    a['inst'] += f(r, 'smsp__inst_executed.sum') or 0.0
    a['regs'] = int(f(r, 'launch__registers_per_thread') or 0)
    a['grid'] = int(f(r, 'launch__grid_size') or 0)
out = {'source': sys.argv[3], 'workload': sys.argv[2], 'kernels': {}}
if len(sys.argv) > 4:
    out['commit'] = sys.argv[4]
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]['us']):
    n = a['n']
    out['kernels'][k] = {'launches_captured': n, 'dram_bytes_per_launch': (a['rd'] + a['wr']) / n, 'dram_read': a['rd'] / n,
                         'dram_write': a['wr'] / n, 'ncu_us': a['us'] / n, 'issue_active_pct': a['issue'] / n,
                         'warp_inst': a['inst'] / n, 'regs': a['regs'], 'grid': a['grid']}
json.dump(out, sys.stdout, indent=1)
