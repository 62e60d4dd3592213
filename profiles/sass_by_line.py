"""Joins an ncu source-page CSV (per-SASS-instruction counts) with `nvdisasm -g` output (line annotations):
executed warp instructions and stall samples per source line.
usage: sass_by_line.py <ncu_source.csv> <nvdisasm_function.sass> <source file> [top]"""
import csv, re, sys, collections
rows = list(csv.reader(open(sys.argv[1]))); hdr = rows[1]; data = rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
A, S, E = ix['Address'], ix['# Samples'], ix['Instructions Executed']
base = int(data[0][A], 16)
cnt = {int(r[A], 16) - base: (int(r[E] or 0), int(r[S] or 0)) for r in data}
line = None; by = collections.defaultdict(lambda: [0, 0, 0])
for l in open(sys.argv[2]):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: line = int(m.group(2)) if m.group(1).endswith(sys.argv[3].split('/')[-1]) else -1; continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/', l)
    if m and line is not None:
        off = int(m.group(1), 16)
        if off in cnt:
            e, s = cnt[off]; by[line][0] += e; by[line][1] += s; by[line][2] += 1
tot_e = sum(v[0] for v in by.values()); tot_s = sum(v[1] for v in by.values())
src = open(sys.argv[3]).read().split('\n')
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
print(f"total executed {tot_e}, samples {tot_s}")
for ln, (e, s, n) in sorted(by.items(), key=lambda kv: -kv[1][0])[:top]:
    text = src[ln - 1].strip()[:110] if 0 < ln <= len(src) else '?'
    print(f"{ln:5d} exec {100*e/tot_e:5.2f}% stall {100*s/tot_s:5.2f}% sass {n:4d} | {text}")
