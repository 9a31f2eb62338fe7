#!/usr/bin/env python
"""configs[4] (256 joints, 49 998 vertices) LBS at one point of the sweep - short enough to sit under ncu.
  python tools/cfg5_bench.py [--coherent 0|1|2]  (0 random bindings, 1 four nearest bones, 2 bound along the limb) [--n 512] [--pos-only] [--compact -1|0|1] [--tile 0|8|16|32] [--reps 10]"""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from piccolo_b200 import capi, workloads as W

ap = argparse.ArgumentParser()
ap.add_argument("--coherent", type=int, default=1)
ap.add_argument("--n", type=int, default=512)
ap.add_argument("--pos-only", action="store_true")
ap.add_argument("--compact", type=int, default=-1)
ap.add_argument("--tile", type=int, default=0)
ap.add_argument("--vsplit", type=int, default=0)
ap.add_argument("--reps", type=int, default=10)
a = ap.parse_args()
peak = 6540.2
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
sk, cl, am, mesh = W.config5_assets({0: False, 1: True, 2: "chain"}[a.coherent])
ctx = capi.Context(0)
ctx.set_option("lbs_compact", a.compact), ctx.set_option("lbs_cta_tile", a.tile), ctx.set_option("lbs_vsplit", a.vsplit)
s5 = ctx.upload_skeleton(sk)
c5 = ctx.upload_clip(s5, cl, am)
m5 = ctx.upload_mesh(mesh, sk.J)
b = ctx.batch(s5, a.n, 1)
b.set(clip_ids=np.full((a.n, 1), c5, np.int32), phase=W.config5_phases(a.n))
b.pose()
fl, ob = (capi.PCL_SKIN_POS, 12) if a.pos_only else (capi.PCL_SKIN_POS_NRM_TAN, 36)
sb = ctx.skinbuf(m5, mesh.V, a.n, fl)
for _ in range(3):
    b.skin(m5, sb)
ctx.sync()
ctx.timer_begin()
for _ in range(a.reps):
    b.skin(m5, sb)
ms = ctx.timer_end() / a.reps
gb = (a.n * (mesh.V * ob + (sk.J + 1) * 48) + mesh.V * 68) / ms / 1e6
print(json.dumps({"coherent": a.coherent, "N": a.n, "out_bytes": ob, "compact": a.compact, "tile": a.tile, "ms": round(ms, 4), "GBps": round(gb, 1),
                  "frac": round(gb / peak, 4), "vertices_per_s": round(a.n * mesh.V / ms * 1e3)}), flush=True)
sb.close(), b.close(), ctx.close()
