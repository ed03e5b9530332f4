"""Summarise the SASS-level stall sampling of one kernel in an .ncu-rep:
   ncu -i X.ncu-rep --page source --csv --launch-skip K --launch-count 1 > sass.csv ; python tools/ncu_stalls.py sass.csv"""
import csv, sys
allr = list(csv.reader(open(sys.argv[1])))
starts = [i for i, x in enumerate(allr) if x and x[0] == 'Kernel Name'] + [len(allr)]
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
r = allr[starts[which]:starts[which + 1]]
print(len(starts) - 1, "kernels in file; showing", which, r[0][1][:100])
h = r[1]; rows = [x for x in r[2:] if len(x) >= len(h) - 2]
col = {name: i for i, name in enumerate(h)}
S = col['# Samples']; I = col['Instructions Executed']
tot = sum(float(x[S] or 0) for x in rows); toti = sum(float(x[I] or 0) for x in rows)
print("samples", tot, "warp-instructions", toti, "sass lines", len(rows))
stalls = [n for n in h if n.startswith('stall_') and 'Not Issued' not in n]
agg = {n: sum(float(x[col[n]] or 0) for x in rows) for n in stalls}
print("stall reasons:", ", ".join("%s %.1f%%" % (n[6:], 100 * v / tot) for n, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
idx = sorted(range(len(rows)), key=lambda k: -float(rows[k][S] or 0))[:N]
for k in sorted(idx):
    x = rows[k]
    top = max(stalls, key=lambda n: float(x[col[n]] or 0))
    print("%5d %6.2f%% inst=%8s %-14s %s" % (k, 100 * float(x[S]) / tot, x[I], top[6:], x[col['Source']].strip()[:90]))
