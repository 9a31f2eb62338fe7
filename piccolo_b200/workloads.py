"""The synthetic crowds of BASELINE.json's configs, defined ONCE so that bench.py and the
full-size parity tests run exactly the same inputs.

Every per-instance input depends only on the GLOBAL instance number, so a rank's shard [first, first + count) of a
sharded run holds the same values as that range of an unsharded run (shard == unsharded).
"""
from __future__ import annotations

import numpy as np

from . import assets as A

CLIP_LEN = 0.76666998863220215  # blend_clip_file_length[0] of player.object.json = 23 frames / 30 fps
CONFIG4_FRAMES = (17, 23, 29, 31, 37, 41, 47)
CONFIG5_SWEEP_N = (1, 8, 64, 512, 4096)


def config2_inputs(first: int, count: int):
    """configs[1]: single clip, phase_i = frac(i * 0.6180339887) in fp32."""
    i = np.arange(first, first + count, dtype=np.float64)
    return np.modf(i * 0.6180339887)[0].astype(np.float32).reshape(count, 1)


def config3_inputs(first: int, count: int):
    """configs[2]: two-clip blend, weights (w, 1 - w) with w_i = frac(i * 0.7548776662); phases a golden-ratio sequence,
    the second clip half a cycle ahead."""
    i = np.arange(first, first + count, dtype=np.float64)
    w = np.modf(i * 0.7548776662)[0].astype(np.float32)
    weight = np.stack([w, np.float32(1) - w], 1)
    phase = np.stack([np.modf(i * 0.6180339887)[0], np.modf(i * 0.6180339887 + 0.5)[0]], 1).astype(np.float32)
    return phase, weight


def config3_clips(cl):
    """The bundled clip + the seeded 17-frame "run" variant (the reference ships one clip)."""
    return [cl, A.make_variant_clip(cl, 17, seed=1234)]


def config4_clips(cl):
    """The bundled clip + 7 seeded variants with frame counts 17..47."""
    return [cl] + [A.make_variant_clip(cl, f, seed=100 + i) for i, f in enumerate(CONFIG4_FRAMES)]


def config4_inputs(first: int, count: int, total: int = 1000000):
    """configs[3]: clip pick = seed-7 stream mod 8, phase = seed-8 uniform [0, 1); generated for the whole crowd and
    sliced, so that every shard sees the unsharded crowd's values."""
    pick = np.random.default_rng(7).integers(0, 8, total).astype(np.int32)
    phase = np.random.default_rng(8).uniform(0, 1, total).astype(np.float32)
    return pick[first:first + count].reshape(count, 1), phase[first:first + count].reshape(count, 1)


def config5_assets(coherent=False):
    """configs[4]: 256-joint balanced tree, 256-channel x 32-frame clip, 49 998-vertex mesh with 4 non-zero weights per
    vertex.  coherent = False: every vertex bound to random bones (no reuse between neighbouring vertices: the worst case);
    coherent = True: tubes around the bones, each vertex bound to its 4 nearest bones (what an exported character looks like)."""
    sk = A.make_synthetic_skeleton(256)
    cl, am = A.make_synthetic_clip(sk, 32)
    if coherent == "chain" or coherent == 2:  # bound along the limb: bone, parent, grandparent, first child
        mesh = A.make_chain_bound_mesh(sk, V=49998, seed=11)
    else:
        mesh = A.make_coherent_mesh(sk, V=49998, seed=11) if coherent else A.make_synthetic_mesh(sk, V=49998)
    return sk, cl, am, mesh


def config5_phases(n: int):
    return np.linspace(0, 1, n, endpoint=False, dtype=np.float32).reshape(n, 1)


def sample_ranges(n_total: int, n_ranges: int, width: int, seed: int = 3):
    """`n_ranges` disjoint [start, start + width) ranges spread over [0, n_total): the first, the last and seeded ones in
    between - the sub-ranges the full-size tests download and hold against the oracle."""
    width = min(width, n_total)
    slots = n_total // width
    n_ranges = min(n_ranges, slots)
    rng = np.random.default_rng(seed)
    picks = {0, slots - 1}
    while len(picks) < n_ranges:
        picks.add(int(rng.integers(0, slots)))
    starts = sorted(p * width for p in picks)
    if len(starts) > 1:
        starts[-1] = n_total - width  # the very last instances, whatever n_total % width is
    return [(s, width) for s in starts]
