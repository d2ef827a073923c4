"""Join an ncu SASS source-page CSV with nvdisasm -g line info: per source line, stall samples and executed instructions.
usage: ncu_lines.py <sass_csv> <nvdisasm_file> <mangled-substring> <source.cu> [top]"""
import csv, re, sys
csv_path, dis_path, sym, src_path = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
# offset -> line
off2line, cur, inside = {}, None, False
for ln in open(dis_path):
    if ln.startswith('.text.') and ln.rstrip().endswith(':'):
        inside = sym in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/', ln)
    if m and cur:
        off2line[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(csv_path)))
hdr = rows[1]
ia, iss, ie, it = hdr.index('Address'), hdr.index('# Samples'), hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed')
base = None
agg = {}
for r in rows[2:]:
    try:
        a = int(r[ia], 16)
    except Exception:
        continue
    if base is None:
        base = a
    line = off2line.get(a - base, ('?', 0))
    s, e, t = int(r[iss]), int(r[ie]), int(r[it])
    x = agg.setdefault(line, [0, 0, 0])
    x[0] += s; x[1] += e; x[2] += t
ts, te = sum(v[0] for v in agg.values()), sum(v[1] for v in agg.values())
src = open(src_path).read().split('\n')
print('samples', ts, 'warp-inst', te)
for line, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    text = src[line[1] - 1].strip()[:100] if line[0] == src_path.split('/')[-1] and line[1] else ''
    print(f"{100*v[0]/ts:5.1f}% smp {100*v[1]/te:5.1f}% inst  act {v[2]/max(v[1],1):4.1f}  {line[0]}:{line[1]}: {text}")
