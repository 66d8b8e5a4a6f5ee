#!/usr/bin/env python3
"""Micro-sweeps of BASELINE.json configs[4] and the secondary shapes (configs[2], the recursive/FRI widths), one GPU.

  * single-column (and 3-, 6-, 18-column) NTT / INTT, 2^20 .. 2^27 rows, against the HBM roofline:
    algorithmic bytes = 16 B per element per transform (read once, write once);
  * linear_hash over 2^24 leaves x {12, 16, 64, 128, 256, 660} columns: permutations/s and GB/s;
  * commits: configs[2] (2^20 x 200, blowup 8), recursive shapes (2^17 x 18 x8, 2^20 x 18 x4), cm4-like 2^23 x 6 x2.
  * widened rows: step-4 quotient commit 2^23 -> 2^24 (qDim 3, qDeg 2), FRI folds 2^24 -> 2^19 -> 2^14 with their
    layer trees, constant tree 2^20 x 100 (blowup 2); device-resident buffers.
  * data independence: config B/2 (2^23 x 330, blowup 2) on uniform inputs and on a "trace-like" distribution
    (50 % zeros, ~30 % below 2^16, ~20 % uniform).
Timing: CUDA events on the launching stream, 2 warm-ups, best of 5; inputs are larger than L2 from 2^24 elements up
(smaller cases are flagged "fits L2").  Prints one JSON object per line."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from xlayer_prover_b200 import Context, tree_num_elements  # noqa: E402

HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
dev = torch.device("cuda", 0)
ctx = Context(0)
stream = torch.cuda.Stream(device=dev)
ctx.set_stream(stream.cuda_stream)
only = set(sys.argv[1:])


def timed(fn, reps=5, warm=2):
    with torch.cuda.stream(stream):
        for _ in range(warm):
            fn()
        best = 1e30
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            fn()
            b.record(stream)
            b.synchronize()
            best = min(best, a.elapsed_time(b))
    return best


def emit(**kw):
    print(json.dumps(kw), flush=True)


if not only or "ntt" in only:
    for ncols in (1, 3, 6, 18):
        for logn in range(20, 28):
            n = 1 << logn
            if n * ncols * 8 * 2 > 100e9:
                continue
            buf = torch.empty(n * ncols, dtype=torch.int64, device=dev)
            out = torch.empty(n * ncols, dtype=torch.int64, device=dev)
            ctx.synth_fill(buf, n, ncols, 0x1234)
            for inverse in (False, True):
                ms = timed(lambda: ctx.ntt(out, buf, n, ncols, inverse, False))
                gbs = 16.0 * n * ncols / ms / 1e6
                emit(kind="ntt", inverse=inverse, log_rows=logn, ncols=ncols, ms=round(ms, 4), algorithmic_GBps=round(gbs, 1),
                     frac_of_measured_hbm=round(gbs / HBM, 4), fits_l2=bool(n * ncols * 16 < 126e6),
                     passes=(logn + 7) // 8, gelem_per_s=round(n * ncols / ms / 1e6, 3))
            del buf, out
            torch.cuda.empty_cache()

if not only or "hash" in only:
    rows = 1 << 24
    for ncols in (12, 16, 64, 128, 256, 660):
        src = torch.empty(rows * ncols, dtype=torch.int64, device=dev)
        out = torch.empty(rows * 4, dtype=torch.int64, device=dev)
        ctx.synth_fill(src, rows, ncols, 0x5678)
        ms = timed(lambda: ctx.linear_hash_rows(out, src, ncols, ncols, rows), reps=3, warm=1)
        perms = rows * ((ncols + 7) // 8)
        emit(kind="linear_hash", rows_log2=24, ncols=ncols, ms=round(ms, 3), perm_per_s=round(perms / ms * 1e3), GBps=round((8.0 * ncols + 32) * rows / ms / 1e6, 1),
             frac_of_measured_hbm=round((8.0 * ncols + 32) * rows / ms / 1e6 / HBM, 4))
        del src, out
        torch.cuda.empty_cache()

if not only or "commit" in only or "stages" in only:
    for name, logn, ncols, logb in (("configs[2] recursive/aggregation shape", 20, 200, 3), ("recursive1 cm1", 17, 18, 3), ("c12a cm1", 20, 18, 2),
                                    ("zkevm cm4 (quotient)", 23, 6, 1), ("zkevm cm2", 23, 168, 1), ("zkevm cm3", 23, 408, 1),
                                    ("zkevm cm1 (fork 9: 751 columns)", 23, 751, 1)):
        if only and "stages" in only and not name.startswith("zkevm"):
            continue
        n, ne = 1 << logn, 1 << (logn + logb)
        src = torch.empty(n * ncols, dtype=torch.int64, device=dev)
        ext = torch.empty(ne * ncols, dtype=torch.int64, device=dev)
        tree = torch.empty(tree_num_elements(ne), dtype=torch.int64, device=dev)
        ctx.synth_fill(src, n, ncols, 0x9E3779B97F4A7C15 + 3)
        ctx.get_profile()
        ctx.set_profiling(True)
        ms = timed(lambda: ctx.commit(ext, tree, src, n, ne, ncols), reps=3, warm=1)
        ctx.set_profiling(False)
        prof = ctx.get_profile()
        k = max(1, prof["leaf_hash"][1])
        emit(kind="commit", shape=name, log_rows=logn, ncols=ncols, blowup=1 << logb, ms=round(ms, 3), melem_per_s=round(n * ncols / ms / 1e3, 1),
             stage_ms={s: round(v[0] / k, 3) for s, v in prof.items()},
             lde_algorithmic_GBps=round(8.0 * ncols * (n + ne) / ((prof["intt"][0] + prof["ntt_ext"][0]) / k) / 1e6, 1))
        del src, ext, tree
        torch.cuda.empty_cache()
if not only or "dist" in only:
    n, ne, ncols = 1 << 23, 1 << 24, 330
    src = torch.empty(n * ncols, dtype=torch.int64, device=dev)
    ext = torch.empty(ne * ncols, dtype=torch.int64, device=dev)
    tree = torch.empty(tree_num_elements(ne), dtype=torch.int64, device=dev)
    roots = {}
    for dist_name in ("uniform", "trace-like"):
        ctx.synth_fill(src, n, ncols, 0x9E3779B97F4A7C15 + 2)
        if dist_name == "trace-like":
            step = 1 << 26
            for o in range(0, src.numel(), step):               # in place, chunked: selector = 4 hashed bits of the value
                v = src[o:o + step]
                sel = ((v * -7046029254386353131) >> 60) & 15
                v.copy_(torch.where(sel < 8, torch.zeros_like(v), torch.where(sel < 13, v & 0xFFFF, v)))
            torch.cuda.synchronize()
        zeros = float((src[: 1 << 24] == 0).float().mean().item())
        ctx.get_profile()
        ctx.set_profiling(True)
        ms = timed(lambda: ctx.commit(ext, tree, src, n, ne, ncols), reps=3, warm=1)
        ctx.set_profiling(False)
        prof = ctx.get_profile()
        k = max(1, prof["leaf_hash"][1])
        emit(kind="data_independence", distribution=dist_name, log_rows=23, ncols=ncols, blowup=2, zero_fraction=round(zeros, 3), ms=round(ms, 3),
             stage_ms={s_: round(v_[0] / k, 3) for s_, v_ in prof.items()}, root0=hex(int(tree[-4].item()) & 0xFFFFFFFFFFFFFFFF))
    del src, ext, tree
    torch.cuda.empty_cache()

if not only or "widened" in only:
    n, ne, qdim, qdeg = 1 << 23, 1 << 24, 3, 2
    q = torch.empty(ne * qdim, dtype=torch.int64, device=dev)
    cm4 = torch.empty(ne * qdim * qdeg, dtype=torch.int64, device=dev)
    tree = torch.empty(tree_num_elements(ne), dtype=torch.int64, device=dev)
    ctx.synth_fill(q, ne, qdim, 0xABCD)
    ctx.get_profile()
    ctx.set_profiling(True)
    ms = timed(lambda: ctx.quotient_commit(cm4, tree, q, n, ne, qdim, qdeg), reps=3, warm=1)
    ctx.set_profiling(False)
    prof = ctx.get_profile()
    k = max(1, prof["leaf_hash"][1])
    byts = 8.0 * (ne * qdim + ne * qdim * qdeg)                    # quotient read once, cm4 written once
    emit(kind="quotient_commit", log_n=23, log_n_ext=24, qdim=qdim, qdeg=qdeg, ms=round(ms, 3), stage_ms={s_: round(v[0] / k, 3) for s_, v in prof.items()},
         transforms_algorithmic_GBps=round(byts / ((prof["intt"][0] + prof["ntt_ext"][0]) / k) / 1e6, 1),
         note="stage intt = INTT(2^24 x 3) + chunk scaling, ntt_ext = NTT(2^24 x 6, 2^23 live rows)")
    del q, cm4, tree
    torch.cuda.empty_cache()
    bits, pol = 24, torch.empty((1 << 24) * 3, dtype=torch.int64, device=dev)
    ctx.synth_fill(pol, 1 << 24, 3, 0xF00D)
    shift_inv = pow(7, 0xFFFFFFFF00000001 - 2, 0xFFFFFFFF00000001)
    for nb, nxt in ((19, 14), (14, 9)):
        out = torch.empty((1 << nb) * 3, dtype=torch.int64, device=dev)
        ms = timed(lambda: ctx.fri_fold(out, pol, bits, nb, [3, 5, 7], shift_inv))
        gbs = 24.0 * ((1 << bits) + (1 << nb)) / ms / 1e6
        aux = torch.empty((1 << nb) * 3, dtype=torch.int64, device=dev)
        ltree = torch.empty(tree_num_elements(1 << nxt), dtype=torch.int64, device=dev)
        ms2 = timed(lambda: ctx.fri_group_tree(aux, ltree, out, nb, nxt))
        emit(kind="fri_layer", pol_bits=bits, step_bits=nb, next_bits=nxt, fold_ms=round(ms, 4), fold_GBps=round(gbs, 1), fold_frac_of_measured_hbm=round(gbs / HBM, 4),
             fits_l2=bool(24 * (1 << bits) < 126e6), layer_tree_ms=round(ms2, 4), layer_tree_perms=(1 << nxt) * ((3 * (1 << (nb - nxt)) + 7) // 8) + (1 << nxt) - 1)
        pol, bits = out, nb
        del aux, ltree
    del pol
    torch.cuda.empty_cache()
    n, ne, npols = 1 << 20, 1 << 21, 100
    src = torch.empty(n * npols, dtype=torch.int64, device=dev)
    ct = torch.empty(2 + ne * npols + tree_num_elements(ne), dtype=torch.int64, device=dev)
    ctx.synth_fill(src, n, npols, 0xC0DE)
    ms = timed(lambda: ctx.const_tree(ct, src, n, ne, npols), reps=3, warm=1)
    emit(kind="const_tree", log_n=20, log_n_ext=21, npols=npols, ms=round(ms, 3), melem_per_s=round(n * npols / ms / 1e3, 1))
ctx.close()
