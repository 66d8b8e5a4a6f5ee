#!/usr/bin/env python3
"""e2e (host-pointer) commit at the bench shape under several column-block plans of the streamed pipeline.
usage: e2e_plan_probe.py log_rows cols "NAME:ENV=VAL,ENV=VAL" ...      (plan lists use '+' for ',': GLP_STREAM_PLAN=32+48+72)
Prints ms per commit (3 iterations per plan) and, with GLP_TRACE=1, the pipeline's timeline."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from xlayer_prover_b200 import Context, tree_num_elements  # noqa: E402

log_rows, C = int(sys.argv[1]), int(sys.argv[2])
n, n_ext = 1 << log_rows, 1 << (log_rows + 1)
SEED = 0x9E3779B97F4A7C15 + 2
ctx = Context(0)
hin, hext, htree = ctx.host_alloc(8 * n * C), ctx.host_alloc(8 * n_ext * C), ctx.host_alloc(8 * tree_num_elements(n_ext))
d = ctx.device_alloc(8 * n * C)
ctx.synth_fill(d, n, C, SEED)
ctx.memcpy(hin, d, 8 * n * C)
ctx.device_free(d)
ctx.close()
for spec in sys.argv[3:]:
    name, _, envs = spec.partition(":")
    env = dict(kv.split("=") for kv in envs.split(",") if kv)
    env = {k: v.replace("+", ",") for k, v in env.items()}
    os.environ.update(env)
    c = Context(0)
    for k in env:
        os.environ.pop(k)
    best = 1e9
    for it in range(3):
        t0 = time.perf_counter()
        root = c.commit(hext, htree, hin, n, n_ext, C)
        dt = (time.perf_counter() - t0) * 1e3
        best = min(best, dt)
        print("plan %-10s iter %d  %8.1f ms  root[0] %#x" % (name, it, dt, int(root[0])), flush=True)
    print("plan %-10s best %8.1f ms   %s" % (name, best, env), flush=True)
    c.close()
