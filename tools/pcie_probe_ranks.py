#!/usr/bin/env python3
"""Concurrent host<->device copy rates with 1, 2, 4, ... WORLD_SIZE ranks active (one process per GPU, torchrun).
Every active rank moves `--gib` GiB of pinned host memory to its GPU (H2D), back (D2H), and both at once on two streams;
rank 0 prints the aggregate GB/s per direction for each group size.  Backs the statement that the end-to-end commit at
N > 1 is bounded by what the HOST moves in both directions together, not by the GPUs.
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29700 tools/pcie_probe_ranks.py [--gib 4]"""
import argparse
import os

import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--gib", type=float, default=4.0)
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(args.gib * (1 << 30)) // 8
h_in = torch.empty(n, dtype=torch.int64).pin_memory()
h_out = torch.empty(n, dtype=torch.int64).pin_memory()
h_in.fill_(rank + 1)
d_a = torch.empty(n, dtype=torch.int64, device="cuda")
d_b = torch.ones(n, dtype=torch.int64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
flag = torch.zeros(1, device="cuda")


def timed(fn, active):
    """max over the active ranks of the device time of `reps` calls of fn, all ranks starting together"""
    best = None
    for _ in range(2):                                    # first round warms up
        dist.all_reduce(flag)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if active:
            for _ in range(args.reps):
                fn()
        for s in (s1, s2):
            torch.cuda.current_stream().wait_stream(s)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) if active else 0.0], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = float(t.item())
    return best


def h2d():
    with torch.cuda.stream(s1):
        d_a.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_b, non_blocking=True)


def both():
    h2d()
    d2h()


g = 1
while g <= world:
    active = rank < g
    gb = g * args.reps * n * 8 / 1e9
    t_up, t_dn, t_both = timed(h2d, active), timed(d2h, active), timed(both, active)
    if rank == 0:
        print("ranks %d   H2D %7.1f GB/s   D2H %7.1f GB/s   both at once %7.1f + %7.1f = %7.1f GB/s" %
              (g, gb / t_up * 1e3, gb / t_dn * 1e3, gb / t_both * 1e3, gb / t_both * 1e3, 2 * gb / t_both * 1e3), flush=True)
    g *= 2
dist.destroy_process_group()
