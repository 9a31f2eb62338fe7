#!/usr/bin/env python
"""Headline benchmark: animated + skinned instances per second (BASELINE.json `metric`).

Workload "config3" (BASELINE.json configs[2]) PER GPU:
  100 000 instances of the bundled 58-bone player skeleton, each blending two clips (the bundled
  crouch-walk clip + a seeded 17-frame synthetic "run" variant -- the reference ships one clip) with
  weights (w, 1-w), then 4-weight linear-blend skinning of a 16 380-vertex synthetic player mesh
  (robot.mesh_bind.json is a missing blob) to positions + normals + tangents (36 B / vertex) into a
  device ring of 25 024 instance slots (14.8 GB; the consumer -- a renderer -- would drain it).
A "step" = one pcl_batch_tick (phase advance + sample + blend + hierarchy + palette) + the LBS launches
that skin all instances (4 ring chunks).  Instances shard across GPUs with no collective (weak scaling).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--instances M]
  torchrun ... bench.py --gpus N ...        (one rank per GPU; rank 0 prints ONE JSON line)

`value`  : device-resident steps, timed with CUDA events on the library's compute stream, max over ranks.
`e2e`    : same step through the C ABI with HOST buffers: per-step host->device upload of every
           instance's phases and weights, and device->host download of every instance's AnimationResult
           (58 joint matrices, rows 0..2 = 48 B each: row 3 is the constant 0 0 0 1 and is filled in by the
           host mirror) into pinned memory, double-buffered (frame i's result lands while frame i + 1 is
           computed; one sync at the end of the timed region); skinned streams stay on the device like the
           reference's vertex-shader output does.  `e2e.host_link_GBps` is the D2H rate this needs from the
           host link over all ranks, `e2e.host_link_ceiling_GBps` the rate of bare pinned copies of the same
           size issued by all ranks at once, measured in the same run.
`roofline`: the LBS kernel (dominant), algorithmic bytes / measured launch time vs the measured HBM peak.
`cpu_baseline` / --impl reference: the reference's own Skeleton code (oracle/_ref) + the mesh.vert
           restatement on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

V_PLAYER = 16380
RING_SLOTS = 25024  # 782 whole 32-instance tiles; 100 000 = 3 x 25 024 + 24 928 (779 tiles)
DT = 1.0 / 60.0
CLIP_LEN = 0.76666998863220215


def measured_hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def build_workload(first, n_inst):
    """Instances [first, first + n_inst) of the global crowd (per-instance inputs depend only on the global id,
    so a rank's shard is identical to the same range of an unsharded run).  piccolo_b200/workloads.py holds the
    definitions; tests/test_gpu_full_size.py checks the same arrays against the oracle."""
    from piccolo_b200 import assets as A, workloads as W
    sk, cl, am, mk = A.load_player()
    run = W.config3_clips(cl)[1]
    mesh = A.make_player_mesh(sk, V=V_PLAYER, seed=42)
    phase, weight = W.config3_inputs(first, n_inst)
    return sk, cl, run, am, mk, mesh, phase, weight


def profiled_traffic():
    """DRAM bytes per LBS launch from the committed ncu capture of this very launch shape (25 000 instances)."""
    p = os.path.join(ROOT, "profiles", "lbs_traffic.json")
    try:
        return json.load(open(p))
    except Exception:
        return None


def lbs_algorithmic_bytes(count, V, J, out_bytes):
    # B_lbs = N*(V*O + (J+1)*P) + V*(36+32); P = 48 B (compact 3x4 palette is what is read)
    return count * (V * out_bytes + (J + 1) * 48) + V * 68


def other_configs(ctx, capi, grp, sk, cl, am, mk, skel, clip0, peak, with_cpu):
    """The other BASELINE.json configs, each beside the reference's CPU figure for the same work when `with_cpu`:
      configs[1]  10 000 instances, single clip, pose only                                  (rank 0)
      configs[3]  1 M-instance crowd, 8 mixed clips, random phases, pose only, SHARDED over all ranks (strong scaling:
                  every rank poses shard_range(1 000 000, world, rank); time = max over ranks, barrier-bracketed)
      configs[4]  256 joints x 49 998 vertices LBS stress: the sweep N in {1, 8, 64, 512, 4096} x {POS, POS+NRM+TAN}
                  on three meshes: random bindings (worst case), every vertex bound to its four NEAREST bones (coherent in
                  space), every vertex bound along its limb (bone, parent, grandparent, child: a painted character)   (rank 0)
    Parity for the same arrays at the same sizes: tests/test_gpu_full_size.py."""
    from piccolo_b200 import workloads as W
    from piccolo_b200.sharding import shard_range
    out = {}
    rank, world = grp.rank, grp.world

    def time_it(fn, reps, warm=3):
        for _ in range(warm):
            fn()
        ctx.sync()
        ctx.timer_begin()
        for _ in range(reps):
            fn()
        return ctx.timer_end() / reps

    cpu_tick = None
    if with_cpu and rank == 0:
        cpu_tick = cpu_tick_rate(sk, cl, am, mk, budget_s=3.0)

    # config 2
    if rank == 0:
        n = 10000
        b = ctx.batch(skel, n, 1)
        b.set(clip_ids=np.full((n, 1), clip0, np.int32), phase=W.config2_inputs(0, n))
        ms = time_it(b.pose, 50)
        gb = n * (12 + 60 * 48) / ms / 1e6
        out["config2_10k_pose"] = {"ms": round(ms, 4), "instances_per_s": round(n / ms * 1e3), "joints_per_s": round(n * sk.J / ms * 1e3),
                                   "GBps": round(gb, 1), "frac": round(gb / peak, 4)}
        if cpu_tick:
            out["config2_10k_pose"]["cpu_baseline"] = cpu_tick
        b.close()
    # config 4, sharded
    total = 1000000
    first, count = shard_range(total, world, rank)
    clips = [clip0] + [ctx.upload_clip(skel, c, am) for c in W.config4_clips(cl)[1:]]
    pick, phase = W.config4_inputs(first, count, total)
    b = ctx.batch(skel, count, 1)
    b.set(clip_ids=np.int32(clips)[pick], phase=phase)
    for _ in range(3):
        b.pose()
    ctx.sync()
    grp.barrier()
    reps = 10
    ctx.timer_begin()
    for _ in range(reps):
        b.pose()
    my_ms = ctx.timer_end() / reps
    grp.barrier()
    ms = grp.max(my_ms)
    b.close()
    if rank == 0:
        gb = total * (12 + 60 * 48) / ms / 1e6
        out["config4_1M_mixed_clips_pose_sharded"] = {
            "ms": round(ms, 4), "rank0_ms": round(my_ms, 4), "n_gpus": world, "instances_per_gpu": count, "scaling": "strong",
            "instances_per_s": round(total / ms * 1e3), "joints_per_s": round(total * sk.J / ms * 1e3),
            "GBps_aggregate": round(gb, 1), "frac_per_gpu": round(gb / world / peak, 4)}
        if cpu_tick:
            out["config4_1M_mixed_clips_pose_sharded"]["cpu_baseline"] = cpu_tick
    # config 5 sweep
    if rank == 0:
        sweep = {}
        for label, coherent in (("random_bindings", False), ("coherent_mesh", True), ("limb_bound_mesh", "chain")):
            sk5, cl5, am5, mesh5 = W.config5_assets(coherent)
            s5 = ctx.upload_skeleton(sk5)
            c5 = ctx.upload_clip(s5, cl5, am5)
            m5 = ctx.upload_mesh(mesh5, sk5.J)
            rows = []
            pal_for_cpu = None
            for n in W.CONFIG5_SWEEP_N:
                b = ctx.batch(s5, n, 1)
                b.set(clip_ids=np.full((n, 1), c5, np.int32), phase=W.config5_phases(n))
                ms_pose = time_it(b.pose, 10)
                if pal_for_cpu is None or len(pal_for_cpu) < 64:
                    m = capi.expand_3x4(b.download_palette3x4(count=min(n, 64)))
                    pal_for_cpu = np.concatenate([np.tile(np.eye(4, dtype=np.float32).reshape(1, 1, 16), (len(m), 1, 1)), m], 1)
                for fl, name, ob in ((capi.PCL_SKIN_POS, "POS", 12), (capi.PCL_SKIN_POS_NRM_TAN, "POS+NRM+TAN", 36)):
                    sb5 = ctx.skinbuf(m5, mesh5.V, n, fl)
                    ms = time_it(lambda: b.skin(m5, sb5), 20 if n <= 512 else 10)
                    gb = (n * (mesh5.V * ob + (sk5.J + 1) * 48) + mesh5.V * 68) / ms / 1e6
                    rows.append({"N": n, "out": name, "ms": round(ms, 4), "pose_ms": round(ms_pose, 4), "vertices_per_s": round(n * mesh5.V / ms * 1e3),
                                 "GBps": round(gb, 1), "frac": round(gb / peak, 4)})
                    sb5.close()
                b.close()
            sweep[label] = rows
            if with_cpu:
                sweep[label + "_cpu_baseline"] = cpu_lbs_rate(mesh5, pal_for_cpu, budget_s=3.0)
        out["config5_256j_50kv_lbs_sweep"] = sweep
    return out


def host_link_ceiling(torch, local, nbytes, barrier, max_over_ranks, reps=6):
    """Aggregate GB/s of bare cudaMemcpyAsync device -> pinned host copies of `nbytes`, every rank copying at once
    (barrier on both sides, CUDA events, max over ranks): the ceiling the end-to-end result download is held against."""
    try:
        dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda:%d" % local)
        host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        st = torch.cuda.Stream(device=local)
        with torch.cuda.stream(st):
            host.copy_(dev, non_blocking=True)
            st.synchronize()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(reps):
                host.copy_(dev, non_blocking=True)
            e1.record(st)
            e1.synchronize()
        ms = max_over_ranks(e0.elapsed_time(e1))
        barrier()
        world = int(os.environ.get("WORLD_SIZE", "1"))
        del dev, host
        return round(world * nbytes * reps / (ms * 1e-3) / 1e9, 2)
    except Exception:
        return None


def palette_pool_check(grp, ctx, capi, skel, clip, n_total=200000, steps=5):
    """SURVEY.md 8f.4 with the driver watching (world >= 2, outside every timed region): all ranks pose their shard of one crowd
    into a palette pool on rank 0's GPU - once with the copy engine (pose locally, push), once attached (the pose kernel's
    stores travel over NVLink) - and rank 0 compares the gathered rows, bit for bit, with its own pose of the whole crowd.
    Every rank reaches every barrier whatever happens locally (errors are agreed on with a max-reduction before each step
    that needs a peer), so a failure here costs this entry, not the bench line."""
    from piccolo_b200.sharding import shard_range, open_gathered_pool
    res, state = {"instances": n_total, "world": grp.world}, {"bad": 0.0, "why": ""}

    def guarded(fn):
        if state["bad"]:
            return None
        try:
            return fn()
        except Exception as e:  # noqa: BLE001
            state["bad"], state["why"] = 1.0, "%s: %s" % (type(e).__name__, str(e)[:160])
            return None

    def agree():
        state["bad"] = grp.max(state["bad"])
        return not state["bad"]

    ph = np.modf(np.arange(n_total, dtype=np.float64) * 0.6180339887)[0].astype(np.float32)
    first, count = shard_range(n_total, grp.world, grp.rank)
    b = guarded(lambda: ctx.batch(skel, count, 1))
    guarded(lambda: b.set(clip_ids=np.full((count, 1), clip, np.int32), phase=ph[first:first + count]))
    pool = None
    try:
        pool = open_gathered_pool(grp, ctx, skel, n_total, consumer=0)  # (holds a broadcast: every rank calls it)
    except Exception as e:  # noqa: BLE001
        state["bad"], state["why"] = 1.0, "%s: %s" % (type(e).__name__, str(e)[:160])
    pushed = got = None
    if agree():
        def timed(fn):
            for _ in range(2):
                fn()
            ctx.sync()
            ctx.timer_begin()
            for _ in range(steps):
                fn()
            return ctx.timer_end() / steps
        ms_local = guarded(lambda: timed(b.pose))
        ms_push = guarded(lambda: timed(lambda: (b.pose(), b.push_to_pool(pool, first))))
        guarded(ctx.sync)
        grp.barrier()
        if grp.rank == 0:
            def fetch():
                view = ctx.batch(skel, n_total, 1)
                try:
                    view.attach_pool(pool, 0)
                    view.adopt_palettes()
                    return view.download_palette3x4()
                finally:
                    view.close()
            pushed = guarded(fetch)
        grp.barrier()
        guarded(lambda: b.attach_pool(pool, first))
        ms_att = guarded(lambda: timed(b.pose))
        guarded(ctx.sync)
        grp.barrier()  # every producer's stores have landed
        if grp.rank == 0:
            got = guarded(fetch)

            def whole():
                w = ctx.batch(skel, n_total, 1)
                try:
                    w.set(clip_ids=np.full((n_total, 1), clip, np.int32), phase=ph)
                    w.pose()
                    return w.download_palette3x4()
                finally:
                    w.close()
            ref = guarded(whole)
            if not state["bad"]:
                res["identical"] = bool(np.array_equal(got, ref)) and bool(np.array_equal(pushed, ref))
        grp.barrier()
        ms = [grp.max(x if x is not None else 0.0) for x in (ms_local, ms_push, ms_att)]
        remote = (n_total - count) * 60 * 48 / max(grp.world - 1, 1)  # bytes one producer sends
        if not state["bad"] and all(ms):
            res.update({"pose_ms_local_target": round(ms[0], 4), "pose_plus_push_ms": round(ms[1], 4), "pose_ms_attached": round(ms[2], 4),
                        "attached_GBps_per_producer": round(remote / ms[2] / 1e6, 1), "pushed_GBps_per_producer": round(remote / ms[1] / 1e6, 1)})
    if b is not None:
        try:
            b.close()
        except Exception:  # noqa: BLE001
            pass
    grp.barrier()
    if pool is not None and grp.rank != 0:
        try:
            pool.close()  # unmap before the owner frees
        except Exception:  # noqa: BLE001
            pass
    grp.barrier()
    if pool is not None and grp.rank == 0:
        try:
            pool.close()
        except Exception:  # noqa: BLE001
            pass
    agree()
    if state["bad"]:
        res["error"] = state["why"] or "a peer rank failed"
    return res


def bind_to_gpu_numa_node(local):
    """Pin this rank's host threads (and thereby its pinned-buffer pages) to the CPUs nearest its GPU: with 8 ranks
    downloading 371 MB per step each, cross-socket traffic otherwise halves the end-to-end number."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return 0


def run_ours(args):
    import torch
    from piccolo_b200 import capi
    from piccolo_b200.sharding import RankGroup, shard_range
    grp = RankGroup(backend="nccl", device=torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))
    rank, local, world = grp.rank, grp.local, grp.world
    torch.cuda.set_device(local)
    numa_cpus = bind_to_gpu_numa_node(local) if world > 1 else 0
    barrier, max_over_ranks = grp.barrier, grp.max

    N = args.instances  # per GPU (weak scaling): the global crowd has N * world instances
    first, count = shard_range(N * world, world, rank)
    assert count == N
    sk, cl, run, am, mk, mesh, phase, weight = build_workload(first, N)
    ctx = capi.Context(local)
    skel = ctx.upload_skeleton(sk)
    c0, c1 = ctx.upload_clip(skel, cl, am), ctx.upload_clip(skel, run, am)
    mh = ctx.upload_mesh(mesh, sk.J)
    b = ctx.batch(skel, N, 2)
    b.set(clip_ids=np.tile(np.int32([c0, c1]), (N, 1)), phase=phase, weight=weight, clip_len=np.tile(np.float32([CLIP_LEN, 17.0 / 30.0]), (N, 1)))
    ring = min(RING_SLOTS, N)
    sb = ctx.skinbuf(mh, mesh.V, ring, capi.PCL_SKIN_POS_NRM_TAN)
    chunks = [(f, min(ring, N - f)) for f in range(0, N, ring)]
    if args.lbs_variant is not None:
        ctx.set_option("lbs_variant", args.lbs_variant)
    if args.lbs_warps:
        ctx.set_option("lbs_warps", args.lbs_warps)
    if args.lbs_vsplit:
        ctx.set_option("lbs_vsplit", args.lbs_vsplit)

    def skin_all():
        for f, n in chunks:
            b.skin(mh, sb, first=f, count=n, slot_first=0)

    def step_device():
        b.tick(DT, capi.PCL_ADVANCE_ALL)
        skin_all()

    # Double-buffered frame loop, the way a host engine would drive it: frame i's AnimationResult lands in pinned buffer i % 2
    # while frame i + 1 is computed; the per-frame inputs go through the batch's pinned staging buffer without a host stall.
    # Every step still uploads all of its inputs and downloads every instance's result; the timed region ends with a sync.
    host_pal = [ctx.pinned((N, sk.J, 12), np.float32) for _ in range(2)]
    if args.export_mode:
        ctx.set_option("export_mode", args.export_mode)
    h_phase, h_weight = phase.copy(), weight.copy()
    frame = [0]

    def step_e2e():
        b.set_frame_async(phase=h_phase, weight=h_weight)  # H2D of this step's inputs
        b.pose()
        b.download_palette3x4(out=host_pal[frame[0] & 1], asynchronous=True)  # D2H of AnimationResult (rows 0..2), overlaps the skinning
        skin_all()
        frame[0] += 1

    # ---- device-resident timing -------------------------------------------------------------
    # nvidia-smi needs ~1 s to start reporting: it samples from the warm-up to the end of the kernel-only passes,
    # all of which keep the GPU under the same load as the timed region in the middle.
    clk = ClockSampler(local)
    clk.__enter__()
    t_load = time.perf_counter()
    while time.perf_counter() - t_load < 1.2:  # at least 3 warm-up steps, and long enough for the sampler to come up
        for _ in range(args.warmup):
            step_device()
        ctx.sync()
    l0 = ctx.launch_count()
    barrier()
    ctx.timer_begin()
    for _ in range(args.steps):
        step_device()
    ms = ctx.timer_end()
    barrier()
    launches = ctx.launch_count() - l0
    ms = max_over_ranks(ms)

    # ---- the dominant kernel alone (roofline) ---------------------------------------------------
    ctx.timer_begin()
    for _ in range(args.steps):
        skin_all()
    lbs_ms = ctx.timer_end() / (args.steps * len(chunks))
    lp = ctx.launch_count()
    ctx.timer_begin()
    for _ in range(args.steps):
        b.pose()
    pose_ms = ctx.timer_end() / args.steps
    pose_launches = (ctx.launch_count() - lp) // args.steps  # 4 = the three order_* launches of the frame-coherent order + the pose kernel
    clk.__exit__()
    clocks = clk.summary()
    peak, peak_src = measured_hbm_peak()
    per_launch = np.mean([lbs_algorithmic_bytes(n, mesh.V, sk.J, 36) for _, n in chunks])
    achieved = per_launch / (lbs_ms * 1e-3) / 1e9
    pose_bytes = N * (2 * 12 + 60 * 48)
    roofline = {"bound": "hbm", "kernel": {None: "lbs_cta_kernel<NT=1,TILES=3,T=32>", -1: "lbs_cta_kernel<NT=1,TILES=3,T=32>", 3: "lbs_cta_kernel<NT=1,TILES=3,T=32>", 2: "lbs_inst_kernel<NT=1>"}.get(args.lbs_variant, "lbs_kernel<NT=1>"), "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": int(per_launch), "launch_ms": round(lbs_ms, 4),
                "frac_of_nominal_8000": round(achieved / 8000.0, 4),
                "pose_kernel": {"launch_ms": round(pose_ms, 4), "launches": int(pose_launches), "achieved": round(pose_bytes / (pose_ms * 1e-3) / 1e9, 1),
                                "frac": round(pose_bytes / (pose_ms * 1e-3) / 1e9 / peak, 4),
                                "joints_per_s": round(N * sk.J / (pose_ms * 1e-3), 0)}}

    tr = profiled_traffic()
    if tr and tr.get("instances_per_launch") == chunks[0][1] and tr.get("vertices") == mesh.V:
        from piccolo_b200.build import kernel_sha16
        if tr.get("kernel_sha16") == kernel_sha16():  # a capture of OTHER kernel sources says nothing about this run
            roofline["traffic"] = tr["dram_bytes_per_launch"]
            roofline["traffic_source"] = tr["source"]
        else:
            roofline["traffic_note"] = "profiles/lbs_traffic.json was captured on other kernel sources (hash %s); not reported" % tr.get("kernel_sha16")

    # ---- end to end through the C ABI with host buffers -------------------------------------------
    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    ctx.timer_begin()
    for _ in range(args.steps):
        step_e2e()
    e2e_ms = ctx.timer_end()
    e2e_wall = (time.perf_counter() - t0) * 1e3
    e2e_ms = max_over_ranks(max(e2e_ms, e2e_wall))  # host-side work counts too
    barrier()

    # ---- what the host link can carry: bare pinned D2H copies of the same size, all ranks at once ---------------------
    link = host_link_ceiling(torch, local, int(host_pal[0].nbytes), barrier, max_over_ranks)

    # ---- the other BASELINE.json configs (parity for them at these sizes: tests/test_gpu_full_size.py) ----
    others = None
    if not args.no_other_configs:
        sb.close()  # give the 14.7 GB ring back first
        sb = None
        others = other_configs(ctx, capi, grp, sk, cl, am, mk, skel, c0, peak, with_cpu=(world == 1 and not args.no_cpu_baseline))
        if world >= 2:
            others = others or {}
            others["palette_pool"] = palette_pool_check(grp, ctx, capi, skel, c0)

    total = N * world
    out = {
        "metric": "animated+skinned instances/sec", "value": round(total * args.steps / (ms * 1e-3), 1), "unit": "instances/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "config3: %d instances/GPU x 58-bone player skeleton, 2-clip blend (bundled clip + seeded 17-frame variant), "
                               "4-weight LBS of a %d-vertex synthetic player mesh to POS+NRM+TAN" % (N, mesh.V),
                   "instances_per_gpu": N, "joints": sk.J, "vertices": mesh.V, "clips_per_instance": 2,
                   "ring_slots": ring, "palette_layout": "3x4 fp32, 48 B/slot, pitch 60 slots", "lbs_variant": args.lbs_variant,
                   "l2": "per-step traffic (%.1f GB written) exceeds the 126 MB L2; no flush needed" % (N * mesh.V * 36 / 1e9),
                   "parallelism": "instances sharded, no collective", "host_cpus_bound_per_rank": numa_cpus},
        "clocks": clocks, "gpu_launches": int(launches),
        "e2e": {"value": round(total * args.steps / (e2e_ms * 1e-3), 1), "unit": "instances/s",
                "h2d_bytes_per_step": int(h_phase.nbytes + h_weight.nbytes), "d2h_bytes_per_step": int(host_pal[0].nbytes),
                "ms_per_step": round(e2e_ms / args.steps, 4), "result_layout": "rows 0..2 of each joint matrix (48 B/bone); row 3 = 0 0 0 1 is constant",
                "export_mode": args.export_mode,
                "host_link_GBps": round(world * host_pal[0].nbytes * args.steps / (e2e_ms * 1e-3) / 1e9, 2),
                "host_link_ceiling_GBps": link, "frac_of_ceiling": round(world * host_pal[0].nbytes * args.steps / (e2e_ms * 1e-3) / 1e9 / link, 4) if link else None},
        "roofline": roofline,
        "joints_per_s": round(total * sk.J * args.steps / (ms * 1e-3), 0),
        "skinned_vertices_per_s": round(total * mesh.V * args.steps / (ms * 1e-3), 0),
    }
    if others and rank == 0:
        out["other_configs"] = others
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(sk, cl, am, mk, mesh, budget_s=args.cpu_budget)
    if sb is not None:
        sb.close()
    b.close(), ctx.close()
    grp.close()
    if rank == 0:
        print(json.dumps(out), flush=True)


# ------------------------------------------------------------------------------------------------
# the reference's CPU implementation of the path, on the host cores
# ------------------------------------------------------------------------------------------------
def _ref_session(sk, cl, am, mk, td):
    from oracle import pyoracle as O
    urls = O.write_player_json(td, sk, cl, am, mk)
    ref = O.RefLib(td)
    return ref, ref.open(urls["skeleton"], urls["clip"], urls["map"], urls["mask"], CLIP_LEN)


def cpu_reference_rate(sk, cl, am, mk, mesh, budget_s, threads):
    """instances/s of (full AnimationComponent::tick + mesh.vert restatement), all `threads` host threads.
    Returns (rate, description).  Uses oracle/_ref when present (kind 'reference'), else the C port."""
    from oracle import pyoracle as O
    if O.RefLib.available():
        with tempfile.TemporaryDirectory() as td:
            ref, h = _ref_session(sk, cl, am, mk, td)
            # calibrate, then size both legs for ~budget/2 seconds each
            r_pose = ref.bench_pose(h, threads, 200, True)
            per_thread = max(200, int(r_pose * (budget_s / 2) / threads))
            r_pose = ref.bench_pose(h, threads, per_thread, True)
            mats, _ = ref.pose(h, np.linspace(0, 1, 64, dtype=np.float32))
            pal = np.concatenate([np.tile(np.eye(4, dtype=np.float32).reshape(1, 1, 16), (64, 1, 1)), mats], 1)
            r_lbs = ref.bench_lbs(mesh, pal[: 2 * threads], threads)
            n_lbs = max(2 * threads, int(r_lbs * (budget_s / 2)))
            pal_n = pal[np.arange(n_lbs) % 64]
            r_lbs = ref.bench_lbs(mesh, pal_n, threads)
        rate = 1.0 / (1.0 / r_pose + 1.0 / r_lbs)
        return rate, "reference", ("oracle/_ref: %d full ticks (getBlendStateWithClipData+applyAnimation+outputAnimationResult; the reference plays "
                                   "clip 0 only, skeleton.cpp:46) at %.0f/s + mesh.vert restatement of %d instances x %d vertices at %.1f/s"
                                   % (per_thread * threads, r_pose, n_lbs, mesh.V, r_lbs))
    orc = O.Oracle()
    sv, cv = orc.skeleton(sk), orc.clip(cl, am)
    n = 20000
    ph = np.random.default_rng(0).uniform(0, 1, n).astype(np.float32)
    t0 = time.perf_counter()
    mats = orc.pose_batch(sv, [cv], np.zeros((n, 1), np.int32), ph, n_threads=threads)
    r_pose = n / (time.perf_counter() - t0)
    t0 = time.perf_counter()
    k = 8
    for i in range(k):
        orc.lbs(mesh, orc.palette_slots(mats[i]))
    r_lbs = k / (time.perf_counter() - t0)
    rate = 1.0 / (1.0 / r_pose + 1.0 / r_lbs)
    return rate, "port", "oracle/oracle.c: %d poses on %d threads + %d single-thread LBS passes" % (n, threads, k)


def cpu_tick_rate(sk, cl, am, mk, budget_s=3.0):
    """The reference's full AnimationComponent::tick of the player skeleton (what configs[1] and configs[3] compute per
    instance; the reference samples one clip whatever the crowd mixes), all host threads, a bounded sample."""
    from oracle import pyoracle as O
    threads = os.cpu_count() or 1
    try:
        if not O.RefLib.available():
            return None
        with tempfile.TemporaryDirectory() as td:
            ref, h = _ref_session(sk, cl, am, mk, td)
            r = ref.bench_pose(h, threads, 200, True)
            per_thread = max(200, int(r * budget_s / threads))
            r = ref.bench_pose(h, threads, per_thread, True)
        return {"value": round(r, 1), "unit": "instances/s", "cores": threads, "kind": "reference",
                "sample": "oracle/_ref: %d full ticks of the 58-bone player skeleton" % (per_thread * threads)}
    except Exception as e:  # a baseline that cannot run must not take the bench line down
        return {"error": str(e)[:200]}


def cpu_lbs_rate(mesh, palettes, budget_s=3.0):
    """mesh.vert restatement (reference matrix operators) of `mesh` on all host threads, POS+NRM+TAN, bounded sample."""
    from oracle import pyoracle as O
    threads = os.cpu_count() or 1
    try:
        if not O.RefLib.available():
            return None
        with tempfile.TemporaryDirectory() as td:
            ref = O.RefLib(td)
            r = ref.bench_lbs(mesh, palettes[np.arange(threads) % len(palettes)], threads)
            n = max(threads, int(r * budget_s))
            r = ref.bench_lbs(mesh, palettes[np.arange(n) % len(palettes)], threads)
        return {"value": round(r, 2), "unit": "instances/s", "vertices_per_s": round(r * mesh.V), "cores": threads, "kind": "reference",
                "sample": "oracle/_ref mesh.vert restatement: %d instances x %d vertices" % (n, mesh.V)}
    except Exception as e:
        return {"error": str(e)[:200]}


def cpu_baseline(sk, cl, am, mk, mesh, budget_s=16.0):
    threads = os.cpu_count() or 1
    rate, kind, sample = cpu_reference_rate(sk, cl, am, mk, mesh, budget_s, threads)
    out = {"value": round(rate, 2), "unit": "instances/s", "cores": threads, "kind": kind, "sample": sample}
    try:  # single-core figures of the reference's own code, 10 000 iterations after warm-up
        from oracle import pyoracle as O
        if O.RefLib.available():
            with tempfile.TemporaryDirectory() as td:
                ref, h = _ref_session(sk, cl, am, mk, td)
                ref.bench_pose(h, 1, 1000, True)
                out["single_core_us_per_instance"] = {"pose_only": round(1e6 / ref.bench_pose(h, 1, 10000, False), 2),
                                                      "full_tick": round(1e6 / ref.bench_pose(h, 1, 10000, True), 2)}
    except Exception:
        pass
    try:
        out["host_cpu"] = [l.split(":", 1)[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
    except Exception:
        pass
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sk, cl, run, am, mk, mesh, phase, weight = build_workload(0, 1024)
    threads = os.cpu_count() or 1
    budget = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    rates, kind, sample = [], "port", ""
    t_all = time.perf_counter()
    for s in range(args.warmup + args.steps):
        r, kind, sample = cpu_reference_rate(sk, cl, am, mk, mesh, budget, threads)
        if s >= args.warmup:
            rates.append(r)
    wall = time.perf_counter() - t_all
    v = float(np.mean(rates))
    out = {"impl": "reference", "metric": "animated+skinned instances/sec", "value": round(v, 2), "unit": "instances/s",
           "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": round(wall * 1e3 / max(1, args.steps + args.warmup), 2), "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": "config3 on the host cores: each step = a bounded sample of the same workload (%s)" % sample},
           "cpu_baseline": {"value": round(v, 2), "unit": "instances/s", "cores": threads, "kind": kind, "sample": sample},
           "e2e": {"value": round(v, 2), "unit": "instances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--instances", type=int, default=100000, help="instances per GPU")
    ap.add_argument("--lbs-variant", type=int, default=None)
    ap.add_argument("--lbs-warps", type=int, default=0)
    ap.add_argument("--lbs-vsplit", type=int, default=0)
    ap.add_argument("--export-mode", type=int, default=0, help="pcl_set_option export_mode for the e2e result download")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true")
    ap.add_argument("--cpu-budget", type=float, default=16.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
