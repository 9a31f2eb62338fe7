#!/usr/bin/env python
"""stdin: bench.py output; prints the last JSON line's key numbers prefixed by argv[1:]."""
import json, sys
line = [l for l in sys.stdin.read().splitlines() if l.startswith("{")]
if not line:
    print(" ".join(sys.argv[1:]), "NO JSON LINE")
    sys.exit(0)
d = json.loads(line[-1])
r = d.get("roofline", {})
print(" ".join(sys.argv[1:]), "value=%.4g" % d["value"], "ms/step=%s" % d["ms_per_step"], "lbs GB/s=%s frac=%s launch_ms=%s" % (r.get("achieved"), r.get("frac"), r.get("launch_ms")),
      "pose_ms=%s" % r.get("pose_kernel", {}).get("launch_ms"), "e2e=%.4g" % d["e2e"]["value"], "clocks=%s" % d.get("clocks"))
