"""The N > 1 path on CPU: two gloo ranks shard a crowd, time their shard and reduce like bench.py does."""
import os
import socket
import subprocess
import sys
import textwrap

import numpy as np

from conftest import ROOT
from piccolo_b200.sharding import shard_range


def test_shard_ranges_tile_the_crowd_exactly():
    for n in (1, 7, 8, 100000, 1000003):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, world, r) for r in range(world)]
            covered = np.zeros(n, np.int32)
            for first, count in spans:
                assert 0 <= first <= n and count >= 0
                covered[first:first + count] += 1
            assert np.all(covered == 1)
            per = -(-n // world)
            assert all(c <= per for _, c in spans)


WORKER = textwrap.dedent("""
    import json, os, sys, time
    sys.path.insert(0, %r)
    import numpy as np
    from piccolo_b200.sharding import RankGroup, shard_range
    from piccolo_b200 import assets as A
    from oracle import pyoracle as O
    g = RankGroup(backend="gloo")
    N = 1001
    first, count = shard_range(N, g.world, g.rank)
    sk, cl, am, mk = A.load_player()
    orc = O.Oracle()
    ph = np.modf(np.arange(N, dtype=np.float64) * 0.6180339887)[0].astype(np.float32)
    g.barrier()
    t0 = time.perf_counter()
    mats = orc.pose_batch(orc.skeleton(sk), [orc.clip(cl, am)], np.zeros((count, 1), np.int32), ph[first:first + count], n_threads=1)
    elapsed = time.perf_counter() - t0 + 0.01 * g.rank          # rank 1 is made the slow one
    g.barrier()
    worst, total = g.max(elapsed), g.sum(float(count))
    handle = g.broadcast_bytes(bytes(range(64)) if g.rank == 0 else bytes(64), 64, src=0)   # how a palette pool's IPC handle travels
    assert handle == bytes(range(64))
    np.save(os.path.join(%r, "shard%%d.npy" %% g.rank), mats)
    if g.rank == 0:
        print(json.dumps({"world": g.world, "total": total, "worst": worst, "mine": elapsed}))
    g.close()
""")


def test_two_gloo_ranks_shard_time_and_reduce(tmp_path, oracle, player):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % (ROOT, str(tmp_path)))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert r.returncode == 0, r.stderr[-2000:]
    import json
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    assert out["world"] == 2 and out["total"] == 1001
    assert out["worst"] >= out["mine"] and out["worst"] >= 0.01  # max over ranks picked the slow rank
    # concatenated shards == the unsharded result, bit for bit
    sk, cl, am, mk = player
    ph = np.modf(np.arange(1001, dtype=np.float64) * 0.6180339887)[0].astype(np.float32)
    whole = oracle.pose_batch(oracle.skeleton(sk), [oracle.clip(cl, am)], np.zeros((1001, 1), np.int32), ph)
    parts = np.concatenate([np.load(tmp_path / "shard0.npy"), np.load(tmp_path / "shard1.npy")])
    assert np.array_equal(parts, whole)
