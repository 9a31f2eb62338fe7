#!/usr/bin/env python
"""Pose-only throughput (configs 2 and 4 of BASELINE.json): K1-K4 fused kernel, device resident."""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from piccolo_b200 import assets as A, capi

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, nargs="+", default=[10000, 100000, 1000000])
ap.add_argument("--k", type=int, nargs="+", default=[1, 2])
ap.add_argument("--clips", type=int, default=8)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--wpb", type=int, default=0)
ap.add_argument("--pose-sort", type=int, default=-1, help="-1 auto, 0 off, 1 on: frame-coherent visiting order")
ap.add_argument("--min-blocks", type=int, default=0)
ap.add_argument("--level", type=int, default=-1, help="-1 auto, 0 off, 1 on: level-parallel kernel for small crowds")
ap.add_argument("--split", type=int, default=-1, help="-1 auto, 0 off, 1 on: two warps per 32 instances, half a skeleton each")
ap.add_argument("--sorted", type=int, default=0, help="1: instances grouped by clip id (what a crowd manager would do)")
args = ap.parse_args()
sk, cl, am, mk = A.load_player()
frames = (17, 23, 29, 31, 37, 41, 47)
clips = [cl] + [A.make_variant_clip(cl, f, seed=100 + i) for i, f in enumerate(frames)][: args.clips - 1]
ctx = capi.Context(0)
ctx.set_option("pose_warps_per_block", args.wpb)
ctx.set_option("pose_sort", args.pose_sort)
ctx.set_option("pose_min_blocks", args.min_blocks)
ctx.set_option("pose_level", args.level)
ctx.set_option("pose_split", args.split)
skel = ctx.upload_skeleton(sk)
ids = np.asarray([ctx.upload_clip(skel, c, am) for c in clips], np.int32)
peak = 6540.2
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
for K in args.k:
    for N in args.n:
        rng = np.random.default_rng(7)
        b = ctx.batch(skel, N, K)
        pick = rng.integers(0, len(ids), (N, K))
        if args.sorted:
            pick = pick[np.argsort(pick[:, 0], kind='stable')]
        b.set(clip_ids=ids[pick], phase=rng.uniform(0, 1, (N, K)).astype(np.float32),
              weight=rng.uniform(0.1, 1, (N, K)).astype(np.float32))
        for _ in range(3):
            b.pose()
        ctx.sync()
        ctx.timer_begin()
        for _ in range(args.steps):
            b.pose()
        ms = ctx.timer_end() / args.steps
        byts = N * (K * 12 + 60 * 48)
        print(json.dumps({"K": K, "N": N, "clips": len(clips), "ms": round(ms, 4), "instances_per_s": round(N / ms * 1e3), "joints_per_s": round(N * 58 / ms * 1e3),
                          "GBps": round(byts / ms / 1e6, 1), "frac_of_measured_hbm": round(byts / ms / 1e6 / peak, 4)}), flush=True)
        b.close()
ctx.close()
