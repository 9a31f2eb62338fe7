#!/usr/bin/env python
"""SURVEY.md 8f.4 on real hardware: every rank poses its shard of a crowd STRAIGHT INTO a palette pool that lives on rank 0's
GPU (CUDA IPC mapping; the pose kernel's palette stores travel over NVLink), rank 0 adopts the rows and compares them with its
own pose of the whole crowd, bit for bit.  Also times the pose with the remote target against the local one.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/pool_gather_check.py
"""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from piccolo_b200 import assets as A, capi
from piccolo_b200.sharding import RankGroup, shard_range, open_gathered_pool

ap = argparse.ArgumentParser()
ap.add_argument("--instances", type=int, default=200000, help="whole crowd (all ranks)")
ap.add_argument("--steps", type=int, default=10)
args = ap.parse_args()

g = RankGroup(backend="nccl")
ctx = capi.Context(g.local)
sk, cl, am, mk = A.load_player()
skel = ctx.upload_skeleton(sk)
clip = ctx.upload_clip(skel, cl, am)
N = args.instances
ph = np.modf(np.arange(N, dtype=np.float64) * 0.6180339887)[0].astype(np.float32)
first, count = shard_range(N, g.world, g.rank)

# local target first: the baseline time
b = ctx.batch(skel, count, 1)
b.set(clip_ids=np.full((count, 1), clip, np.int32), phase=ph[first:first + count])
for _ in range(3):
    b.pose()
ctx.sync()
g.barrier()
ctx.timer_begin()
for _ in range(args.steps):
    b.pose()
ms_local = g.max(ctx.timer_end() / args.steps)

# pose into the local buffer, then push it to rank 0's pool with the copy engine
pool = open_gathered_pool(g, ctx, skel, N, consumer=0)
for _ in range(2):
    b.pose(), b.push_to_pool(pool, first)
ctx.sync()
g.barrier()
ctx.timer_begin()
for _ in range(args.steps):
    b.pose(), b.push_to_pool(pool, first)
ms_push = g.max(ctx.timer_end() / args.steps)
ctx.sync()
g.barrier()
ok_push = True
if g.rank == 0:
    view = ctx.batch(skel, N, 1)
    view.attach_pool(pool, 0)
    view.adopt_palettes()
    pushed = view.download_palette()
    view.close()

# now the shard's palettes live in rank 0's pool
b.attach_pool(pool, first)
for _ in range(3):
    b.pose()
ctx.sync()
g.barrier()
ctx.timer_begin()
for _ in range(args.steps):
    b.pose()
ms_pool = g.max(ctx.timer_end() / args.steps)
ctx.sync()
g.barrier()  # every producer's stores have landed

ok = True
if g.rank == 0:
    view = ctx.batch(skel, N, 1)
    view.attach_pool(pool, 0)
    view.adopt_palettes()
    got = view.download_palette()
    view.close()
    whole = ctx.batch(skel, N, 1)
    whole.set(clip_ids=np.full((N, 1), clip, np.int32), phase=ph)
    whole.pose()
    ref = whole.download_palette()
    whole.close()
    ok = bool(np.array_equal(got, ref)) and bool(np.array_equal(pushed, ref))
    remote_bytes = (N - count) * 60 * 48
    print(json.dumps({"world": g.world, "instances": N, "identical": ok, "pose_ms_local_target": round(ms_local, 4),
                      "pose_ms_pool_target": round(ms_pool, 4), "pose_plus_push_ms": round(ms_push, 4), "remote_palette_bytes_per_pose": remote_bytes,
                      "remote_GBps_if_pose_bound_by_link": round(remote_bytes / max(g.world - 1, 1) / ms_pool / 1e6, 1)}), flush=True)
g.barrier()
b.close()
if g.rank != 0:
    pool.close()  # unmap before the owner frees
g.barrier()
if g.rank == 0:
    pool.close()
ctx.close()
g.close()
sys.exit(0 if ok else 1)
