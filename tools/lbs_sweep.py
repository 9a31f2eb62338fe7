#!/usr/bin/env python
"""Config 5 of BASELINE.json: synthetic 256-joint skeleton, 49 998-vertex mesh, 4 non-zero weights per vertex;
sweep N and the output flags, report achieved algorithmic GB/s of the LBS launch (SURVEY.md 8d formula)."""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from piccolo_b200 import assets as A, capi

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, nargs="+", default=[1, 8, 64, 512, 4096])
ap.add_argument("--joints", type=int, default=256)
ap.add_argument("--vertices", type=int, default=49998)
ap.add_argument("--coherent", type=int, default=0, help="1: tube mesh with neighbouring vertices sharing bones; 0: SURVEY's random binding")
ap.add_argument("--variant", type=int, default=-1)
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--inst-tile", type=int, default=0)
ap.add_argument("--vsplit", type=int, default=0)
ap.add_argument("--warps", type=int, default=0)
ap.add_argument("--cta-tile", type=int, default=0)
args = ap.parse_args()
peak = 6540.2
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
sk = A.make_synthetic_skeleton(args.joints)
cl, am = A.make_synthetic_clip(sk, 32)
mesh = A.make_player_mesh(sk, V=args.vertices, seed=9) if args.coherent else A.make_synthetic_mesh(sk, V=args.vertices)
ctx = capi.Context(0)
skel = ctx.upload_skeleton(sk)
clip = ctx.upload_clip(skel, cl, am)
mh = ctx.upload_mesh(mesh, sk.J)
ctx.set_option("lbs_variant", args.variant)
ctx.set_option("lbs_inst_tile", args.inst_tile)
ctx.set_option("lbs_vsplit", args.vsplit)
ctx.set_option("lbs_warps", args.warps)
ctx.set_option("lbs_cta_tile", args.cta_tile)
for flags, O in ((capi.PCL_SKIN_POS, 12), (capi.PCL_SKIN_POS_NRM_TAN, 36)):
    for N in args.n:
        b = ctx.batch(skel, N, 1)
        b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=np.linspace(0, 1, N, endpoint=False, dtype=np.float32))
        b.pose()
        sb = ctx.skinbuf(mh, mesh.V, N, flags)
        for _ in range(3):
            b.skin(mh, sb)
        ctx.sync()
        ctx.timer_begin()
        for _ in range(args.steps):
            b.skin(mh, sb)
        ms = ctx.timer_end() / args.steps
        byts = N * (mesh.V * O + (sk.J + 1) * 48) + mesh.V * 68
        print(json.dumps({"J": sk.J, "V": mesh.V, "N": N, "out": "POS" if O == 12 else "POS+NRM+TAN", "ms": round(ms, 4), "GBps": round(byts / ms / 1e6, 1),
                          "frac_of_measured_hbm": round(byts / ms / 1e6 / peak, 4), "vertices_per_s": round(N * mesh.V / ms * 1e3)}), flush=True)
        sb.close(), b.close()
ctx.close()
