#!/usr/bin/env python
"""Wall-clock latency of ONE AnimationComponent::tick equivalent through the C ABI (batch of N, default 1): device-side phase
advance + pose, then the synchronous AnimationResult download - what the host mirror's AnimationComponent::tick does."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from piccolo_b200 import assets as A, capi

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, nargs="+", default=[1, 32, 1000])
ap.add_argument("--iters", type=int, default=2000)
ap.add_argument("--level", type=int, default=-1)
ap.add_argument("--mirror", type=int, nargs="*", default=None,
                help="component counts: per-frame cost through the host C++ mirror (n x AnimationComponent::tick, then n x getResult)")
args = ap.parse_args()
if args.mirror is not None:
    import subprocess, tempfile
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import hostutil as H
    with tempfile.TemporaryDirectory() as td:
        urls = H.write_player_tree(td, A.load_player(), sparse=True)
        for n in (args.mirror or [1, 8, 100]):
            r = H.run(["latency", td, urls["object"], n, args.iters])
            print(r.stdout.strip() or r.stderr.strip(), flush=True)
    sys.exit(0)
sk, cl, am, mk = A.load_player()
ctx = capi.Context(0)
ctx.set_option("pose_level", args.level)
skel = ctx.upload_skeleton(sk)
clip = ctx.upload_clip(skel, cl, am)
for N in args.n:
    b = ctx.batch(skel, N, 1)
    b.set(clip_ids=np.full((N, 1), clip, np.int32), phase=np.linspace(0, 0.9, N, dtype=np.float32),
          clip_len=np.full((N, 1), 0.76666998863220215, np.float32))
    out = ctx.pinned((N, sk.J, 16), np.float32)
    for _ in range(50):
        b.tick(1 / 60.0), b.download_palette(out=out)
    t0 = time.perf_counter()
    for _ in range(args.iters):
        b.tick(1 / 60.0)
        b.download_palette(out=out)
    us = (time.perf_counter() - t0) / args.iters * 1e6
    t0 = time.perf_counter()
    for _ in range(args.iters):
        b.tick(1 / 60.0)
    ctx.sync()
    us_tick = (time.perf_counter() - t0) / args.iters * 1e6
    cl_ids, ph = np.full((N, 1), clip, np.int32), np.linspace(0, 0.9, N, dtype=np.float32).reshape(N, 1)
    for _ in range(50):
        b.frame(clip_ids=cl_ids, phase=ph)
    t0 = time.perf_counter()
    for _ in range(args.iters):
        b.frame(phase=ph)  # pcl_batch_frame: inputs + pose + rows, one launch and one sync
    us_frame = (time.perf_counter() - t0) / args.iters * 1e6
    print(json.dumps({"N": N, "tick_plus_download_us": round(us, 1), "tick_only_back_to_back_us": round(us_tick, 1),
                      "frame_call_us": round(us_frame, 1), "per_instance_us": round(us_frame / N, 3)}), flush=True)
    b.close()
ctx.close()
