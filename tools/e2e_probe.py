#!/usr/bin/env python3
"""e2e (host-pointer) commit at the bench shape for several pipeline block widths; prints ms per commit and,
with GLP_TRACE=1, the pipeline timeline.  usage: e2e_probe.py [log_rows] [cols] width [width ...]
E2E_MODE=full (default: extension + tree to the host) | root (no download at all) | tree (tree only)
        | separate (the reference's unchanged call sequence: extendPol host -> host, then merkelize on the host-resident extension)"""
import ctypes
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from xlayer_prover_b200 import Context, tree_num_elements  # noqa: E402

log_rows, C = int(sys.argv[1]), int(sys.argv[2])
widths = [int(x) for x in sys.argv[3:]]
n, n_ext = 1 << log_rows, 1 << (log_rows + 1)
SEED = 0x9E3779B97F4A7C15 + 2
ctx = Context(0)
hin, hext, htree = ctx.host_alloc(8 * n * C), ctx.host_alloc(8 * n_ext * C), ctx.host_alloc(8 * tree_num_elements(n_ext))
d = ctx.device_alloc(8 * n * C)
ctx.synth_fill(d, n, C, SEED)
ctx.memcpy(hin, d, 8 * n * C)
ctx.device_free(d)
ctx.close()
for w in widths:
    os.environ["GLP_STREAM_COLS"] = str(w)
    c = Context(0)
    roots = []
    for it in range(3):
        t0 = time.perf_counter()
        mode = os.environ.get("E2E_MODE", "full")
        if mode == "separate":
            c.extend_pol(hext, hin, n_ext, n, C)
            t1 = time.perf_counter()
            c.merkletree(htree, hext, C, n_ext)
            t2 = time.perf_counter()
            print("   extendPol %.1f ms, merkelize %.1f ms" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3), flush=True)
            r4 = np.zeros(4, dtype=np.uint64)
            c.memcpy(r4, htree + 8 * (tree_num_elements(n_ext) - 4), 32)
            roots.append(r4)
        else:
            roots.append(c.commit(hext if mode == "full" else None, None if mode == "root" else htree, hin, n, n_ext, C))
        dt = time.perf_counter() - t0
        print("width %4d  iter %d  %8.1f ms  root[0] %#x" % (w, it, dt * 1e3, int(roots[-1][0])), flush=True)
    c.close()
