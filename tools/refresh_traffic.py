#!/usr/bin/env python
"""profiles/lbs_traffic.json from an ncu capture of the bench's LBS launch (read here, no GPU needed).

  python tools/refresh_traffic.py gpurun_out/<capture>.ncu-rep <instances per launch> <vertices> [kernel regex]

The file carries the hash of the kernel sources it was measured on (piccolo_b200.build.kernel_sha16()); bench.py reports
`roofline.traffic` only while that hash matches the sources it runs, so a capture can never outlive the kernel it describes."""
import csv, json, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from piccolo_b200.build import kernel_sha16

rep, n_inst, verts = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
pat = re.compile(sys.argv[4] if len(sys.argv) > 4 else "lbs_cta_kernel")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
H, U = rows[0], rows[1]
hit = [r for r in rows[2:] if pat.search(r[H.index("Kernel Name")])]
if not hit:
    sys.exit("no launch matching %s in %s" % (pat.pattern, rep))
r = hit[-1]


def val(name):
    v, u = float(r[H.index(name)]), U[H.index(name)]
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]


out = {"dram_bytes_per_launch": int(val("dram__bytes_read.sum") + val("dram__bytes_write.sum")), "instances_per_launch": n_inst, "vertices": verts,
       "kernel": r[H.index("Kernel Name")][:80], "ncu_duration_ms": float(r[H.index("gpu__time_duration.sum")]) * {"us": 1e-3, "ms": 1.0, "ns": 1e-6}[U[H.index("gpu__time_duration.sum")]],
       "kernel_sha16": kernel_sha16(),
       "source": "%s (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, one launch of %d instances)" % (os.path.basename(rep), n_inst)}
json.dump(out, open(os.path.join(ROOT, "profiles", "lbs_traffic.json"), "w"), indent=1)
print(json.dumps(out))
