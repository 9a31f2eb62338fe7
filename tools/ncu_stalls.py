#!/usr/bin/env python
"""Where a kernel's warps wait: per-SASS-instruction stall samples from an .ncu-rep source page (read here, no GPU).
  python tools/ncu_stalls.py <rep> [top N] [kernel regex] -> totals per stall reason, then the N instructions with most samples
  (a report that holds several kernels needs the regex: the source pages of all of them would be added up)."""
import csv, subprocess, sys, collections
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
flt = ["--kernel-name", "regex:" + sys.argv[3]] if len(sys.argv) > 3 else []
raw = subprocess.run(["ncu", "-i", rep] + flt + ["--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
H = rows[1]
ci = {h: i for i, h in enumerate(H)}
reasons = [h for h in H if h.startswith('stall_') and 'Not Issued' not in h]
tot = collections.Counter()
lines = []
seen = set()
for n, r in enumerate(rows[2:]):
    if len(r) < len(H) or not r[0].startswith('0x') or r[0] in seen:  # (some reports list a function's rows twice)
        continue
    seen.add(r[0])
    s = int(r[ci['# Samples']] or 0)
    per = {k: int(r[ci[k]] or 0) for k in reasons}
    for k, v in per.items():
        tot[k] += v
    lines.append((s, n, r[ci['Source']].strip(), per, int(r[ci['Instructions Executed']] or 0)))
all_s = sum(t[0] for t in lines)
print("samples", all_s, " instructions executed", sum(t[4] for t in lines))
for k, v in tot.most_common(12):
    print("  %-26s %8d  %.1f%%" % (k, v, 100.0 * v / max(all_s, 1)))
print("top instructions by samples (index, samples, share, main reasons):")
for s, n, src, per, ex in sorted(lines, key=lambda t: -t[0])[:top]:
    main = ', '.join("%s %d" % (k.replace('stall_', ''), v) for k, v in sorted(per.items(), key=lambda kv: -kv[1])[:2] if v)
    print("  #%-5d %6d %5.1f%%  %-70s %s" % (n, s, 100.0 * s / max(all_s, 1), src[:70], main))
