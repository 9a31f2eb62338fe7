#!/usr/bin/env python
"""bench_line.py FILE [label...]: prints the key numbers of the last JSON line of a bench.py output file.
(Takes a file, never stdin: a tool that waits on an empty stdin hangs a GPU box until the call's limit.)"""
import json, os, sys
if len(sys.argv) < 2 or not os.path.exists(sys.argv[1]):
    print("usage: bench_line.py FILE [label...]")
    sys.exit(2)
line = [l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")]
sys.argv = sys.argv[:1] + sys.argv[2:]
if not line:
    print(" ".join(sys.argv[1:]), "NO JSON LINE")
    sys.exit(0)
d = json.loads(line[-1])
r = d.get("roofline", {})
print(" ".join(sys.argv[1:]), "value=%.4g" % d["value"], "ms/step=%s" % d["ms_per_step"], "lbs GB/s=%s frac=%s launch_ms=%s" % (r.get("achieved"), r.get("frac"), r.get("launch_ms")),
      "pose_ms=%s" % r.get("pose_kernel", {}).get("launch_ms"), "e2e=%.4g" % d["e2e"]["value"], "clocks=%s" % d.get("clocks"))
