#!/usr/bin/env python
"""Per-SASS-instruction view of an .ncu-rep source page: shared-memory wavefronts and instruction mix."""
import csv, subprocess, sys, collections, re
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
H = rows[1]
ci = {h: i for i, h in enumerate(H)}
tot_w = 0
mix = collections.Counter()
exec_tot = 0
lines = []
for r in rows[2:]:
    if len(r) < len(H) or not r[0].startswith('0x'):
        continue
    ex = int(r[ci['Instructions Executed']] or 0)
    op = r[ci['Source']].split()[0] if not r[ci['Source']].strip().startswith('@') else r[ci['Source']].split()[1]
    op = op.split('.')[0]
    mix[op] += ex
    exec_tot += ex
    w = int(r[ci['L1 Wavefronts Shared']] or 0)
    wi = int(r[ci['L1 Wavefronts Shared Ideal']] or 0)
    if w:
        tot_w += w
        lines.append((w, wi, ex, r[ci['Source']].strip()))
print("total instr executed", exec_tot, " shared wavefronts", tot_w)
agg = collections.defaultdict(lambda: [0, 0, 0])
for w, wi, ex, s in lines:
    key = re.sub(r'R\d+|UR\d+|0x[0-9a-f]+|\[.*\]', '', s).split()[0:2]
    key = ' '.join(key)
    agg[key][0] += w; agg[key][1] += wi; agg[key][2] += ex
for k, (w, wi, ex) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print("  %-28s wavefronts %12d ideal %12d exec %11d  -> %.2f wf/instr (ideal %.2f)" % (k, w, wi, ex, w / max(ex, 1), wi / max(ex, 1)))
print("instruction mix (top 25):")
for op, n in mix.most_common(25):
    print("  %-10s %12d  %.1f%%" % (op, n, 100.0 * n / exec_tot))
if len(sys.argv) > 2:
    for w, wi, ex, s in sorted(lines, key=lambda t: -t[0])[:int(sys.argv[2])]:
        print("%10d %10d %9d  %s" % (w, wi, ex, s))
