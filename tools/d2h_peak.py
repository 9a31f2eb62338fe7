#!/usr/bin/env python
"""Host-link ceiling: bare device -> pinned-host copies, every rank copying at once.

  python tools/d2h_peak.py [--mb 278]                                  one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/d2h_peak.py

Prints one JSON line (rank 0): aggregate GB/s over all ranks for linear copies of --mb megabytes and for a 2-D copy
of the same payload in 2784-byte rows out of a 2880-byte pitch (the shape of the compact palette read-back).
bench.py measures the linear figure in its own run as `e2e.host_link_ceiling_GBps`.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from bench import host_link_ceiling
    from piccolo_b200.sharding import RankGroup
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=float, default=278.4)
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    grp = RankGroup(backend="nccl", device=torch.device("cuda", local))
    torch.cuda.set_device(local)
    nbytes = int(a.mb * 1e6)
    out = {"n_gpus": grp.world, "bytes_per_rank": nbytes,
           "linear_GBps": host_link_ceiling(torch, local, nbytes, grp.barrier, grp.max),
           "linear_64MB_GBps": host_link_ceiling(torch, local, 64 << 20, grp.barrier, grp.max)}
    # strided rows, as cudaMemcpy2DAsync moves them (torch issues a 2-D copy for a strided source view)
    rows = nbytes // 2784
    dev = torch.empty((rows, 2880), dtype=torch.uint8, device="cuda:%d" % local)
    host = torch.empty((rows, 2784), dtype=torch.uint8, pin_memory=True)
    st = torch.cuda.Stream(device=local)
    with torch.cuda.stream(st):
        host.copy_(dev[:, 48:48 + 2784], non_blocking=True)
        st.synchronize()
        grp.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(6):
            host.copy_(dev[:, 48:48 + 2784], non_blocking=True)
        e1.record(st)
        e1.synchronize()
    ms = grp.max(e0.elapsed_time(e1))
    out["strided_2784_of_2880_GBps"] = round(grp.world * rows * 2784 * 6 / (ms * 1e-3) / 1e9, 2)
    grp.close()
    if grp.rank == 0:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
