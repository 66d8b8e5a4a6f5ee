#!/usr/bin/env python3
"""bench.py -- the STARK-commitment hot path on B200: LDE (extendPol) + Poseidon Merkle commit.

    python bench.py --gpus 1 --steps K --warmup W                 (one GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           bench.py --gpus N --steps K --warmup W                  (N GPUs, one rank each, NCCL)
    python bench.py --impl reference ...                           (the reference's own CPU code, host cores)

Workload (BASELINE.json configs[1]): synthetic 2^23 rows x 660 committed columns, blowup 2 -> 2^24 rows,
extendPol + MerkleTreeGL root.  A "step" is one full commit of that matrix.  The metric is throughput, so
that the CPU arms -- which run a bounded column sample of the same workload -- are directly comparable:
    value [Melem/s] = 10^-6 * (N rows * C columns committed) / seconds per commit      (higher is better)
`commit_ms` and `poseidon_hash_per_s` (BASELINE.json's two raw figures) ride along in the same JSON line.

  value : inputs already resident in HBM, CUDA events on the launching stream, max over ranks.
  e2e   : the same commit through the C ABI with HOST (pinned) buffers: H2D of the trace and D2H of the
          extended matrix + every Merkle node inside the timed region.
  N > 1 : column-sharded LDE whose last NTT pass stores straight into the peers' row slabs over NVLink (the exchange is
          fused into the transform: no all-to-all) -> row-sharded subtrees -> top levels on rank 0
          (xlayer_prover_b200/sharded.py); total work is fixed, so "scaling": "strong".
Only the cpu_baseline leg and --impl reference execute anything under oracle/ (as the thing measured there);
the parity spot-check after the timed region uses it as the checker.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

P = 0xFFFFFFFF00000001
SEED = 0x9E3779B97F4A7C15 + 2          # seed = golden ratio + config id
# One Poseidon permutation, counted two ways.
# (1) ALGORITHMIC: 1122 field multiplies (118 S-boxes x 4, the 12x12 pre-matrix, 22 partial
#     rounds x 23) of four 32x32->64 products each, plus the 7 MDS layers' 12 x 12 small-constant products on both 32-bit halves
#     as DFMA.  Priced at the IMAD.WIDE and DFMA rates glp_measure_pipe_peaks measures in the same run, this is the ceiling
#     `roofline` reports against: it charges nothing for carries, folds or loads.
POS_ALG_IMAD_WIDE = 1122 * 4
POS_ALG_DFMA = 7 * 2 * 144
# (2) The kernel's own instruction stream (tools/sass_leaf.py on the final binary, profiles/r02_sass_mix_leaf.txt): warp
#     instructions per permutation by pipe; costs measured in round 1 (profiles/r01_int_pipes.txt): IMAD.WIDE 4 cycles of the
#     fmaheavy pipe per warp instruction per SM sub-partition, other IMAD, DFMA/I2F and every ALU op 2.  `roofline_sass` is
#     the multiplier-pipe bound of that stream (how close the kernel runs to its own SASS, not to the algorithm).
POS_SASS = {"instructions": 17409, "imad_wide": 4955, "imad_other": 1040, "fp64": 2304, "alu": 7628}
POS_PIPE_CYCLES = {"fmaheavy": 4 * POS_SASS["imad_wide"] + 2 * POS_SASS["imad_other"], "fp64": 2 * POS_SASS["fp64"],
                   "alu": 2 * POS_SASS["alu"], "issue": POS_SASS["instructions"]}
POS_BOUND_CYCLES = POS_PIPE_CYCLES["fmaheavy"] + POS_PIPE_CYCLES["fp64"]
# root of the default workload (2^23 x 660, blowup 2, SEED): the N = 1 run checks it against the reference build through
# whole columns and Merkle paths (verify), every N compares its root with it (verify.root_equals_n1)
KNOWN_ROOTS = {(23, 660, 1): [0x2954300a93d8e1ed, 0xb89e6e7f7a6d3906, 0x8e88252716f12068, 0xc8cda114b642004f]}
# dram__bytes_read.sum + dram__bytes_write.sum of k_linear_hash_rows from the ncu --set full capture at 2^19 rows x 660
# columns (profiles/r01_ncu_full_summary_final2.txt, final binary of round 1: 2.834834 GB + 33.56 MB), per row; a launch's
# traffic scales with its rows
LEAF_DRAM_BYTES_PER_ROW_660 = (2.834834e9 + 33.558016e6) / (1 << 19)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log-rows", type=int, default=23)
    ap.add_argument("--cols", type=int, default=660)
    ap.add_argument("--log-blowup", type=int, default=1)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-e2e-resident", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=25.0, help="target seconds of CPU work for the cpu_baseline sample")
    return ap.parse_args()


def perms_per_commit(n_ext, ncols):
    return n_ext * ((ncols + 7) // 8) + n_ext - 1 if ncols > 4 else n_ext - 1


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# --------------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------------
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        if os.environ.get("BENCH_NO_CLOCK_SAMPLER"):       # diagnostics only (tools/gap_probe.sh): is the sampler itself visible in the timing?
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------------
# CPU arms (the reference's own code on the host cores)
# --------------------------------------------------------------------------------------------
def cpu_oracle():
    from oracle import refbind
    ref = refbind.reference()
    if ref is not None:
        return ref, "reference"
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, capture_output=True)
    return refbind.port(), "port"


CPU_SAMPLE_COLS = 64                     # the CPU arms time C' = 64 of the C columns at the full 2^23 rows


_HOST_CORES = None


def host_cores():
    """CPUs this process may run on, read once: with OMP_PROC_BIND set libgomp pins the main thread to one place when it
    starts, after which sched_getaffinity of the main thread says 1."""
    global _HOST_CORES
    if _HOST_CORES is None:
        try:
            _HOST_CORES = len(os.sched_getaffinity(0))
        except Exception:
            _HOST_CORES = os.cpu_count() or 1
    return _HOST_CORES


def run_cpu_baseline(args, budget_s, nsteps=1, warmup=0):
    """Times the reference's own CPU code (oracle/_ref when it compiled here, else the C port) on a bounded sample of the
    workload and extrapolates to the full job, component by component:
        extendPol      on C' = 64 of the C columns at the full N rows            -> t_lde,    linear in columns
        merkletree     of the first R' rows x C' columns of that extension       -> t_tree   (leaf hashes + levels of R' leaves)
        merkletree     of R' rows x 4 columns (leaves copied: the level loop)    -> t_lev
        t_full = t_lde * C/C' + (t_tree - t_lev) * (N_ext/R') * ceil(C/8)/ceil(C'/8) + t_lev * (N_ext - 1)/(R' - 1)
    The tree levels are NOT scaled by the column count (they do not depend on it), and rows are independent, so both
    scalings are exact in the permutation count; the LDE is linear in columns at >= 32 columns on these hosts."""
    from xlayer_prover_b200._capi import synth_matrix
    orc, kind = cpu_oracle()
    cores = host_cores()
    orc.set_threads(cores)                # torch.distributed.run exports OMP_NUM_THREADS=1; the reference arm uses every core it may
    n, n_ext = 1 << args.log_rows, 1 << (args.log_rows + args.log_blowup)
    C, c = args.cols, min(CPU_SAMPLE_COLS, args.cols)
    a = synth_matrix(n, c, SEED, ncols_total=C)
    # calibrate the leaf sample on 2^14 rows: permutations/s of this host
    cal_rows = min(n_ext, 1 << 14)
    t0 = time.perf_counter()
    orc.merkletree(np.ascontiguousarray(np.resize(a, (cal_rows, c))))
    perm_s = (cal_rows * ((c + 7) // 8) + cal_rows) / max(1e-6, time.perf_counter() - t0)
    per_step = budget_s / max(1, nsteps + warmup)
    lde_est = 0.083 * c * (16.0 / cores) * (n / float(1 << 23))           # round-1 record: 0.083 s per column at 2^23 rows on 16 cores
    leaf_budget = max(0.4, per_step - lde_est)
    leaf_rows = 1 << 12
    while leaf_rows * 2 <= n_ext and (2 * leaf_rows) * ((c + 7) // 8 + 2) / perm_s <= leaf_budget:
        leaf_rows *= 2
    leaf_rows = min(leaf_rows, n_ext)
    steps = []
    for i in range(warmup + nsteps):
        t = orc.timed_parts(a, n_ext, leaf_rows)
        if i >= warmup:
            steps.append(t)
    t_lde, t_tree, t_lev = (sum(x[k] for x in steps) / len(steps) for k in range(3))
    t_leaf = max(0.0, t_tree - t_lev)
    pc, pC = (c + 7) // 8, (C + 7) // 8
    full_lde = t_lde * C / c
    full_leaf = t_leaf * (n_ext / leaf_rows) * (pC / pc if c > 4 else 0)
    full_lev = t_lev * (n_ext - 1) / max(1, leaf_rows - 1)
    t_full = full_lde + full_leaf + full_lev
    return {
        "value": n * C / t_full / 1e6, "unit": "Melem/s", "cores": cores, "kind": kind, "extrapolated": True,
        "sample": "per step: extendPol of 2^%d rows x %d of %d columns (blowup %d), merkletree of the first 2^%d of its 2^%d rows, "
                  "level loop of 2^%d leaves; measured %.2f + %.2f + %.2f s, extrapolated to the full job %.1f + %.1f + %.1f s "
                  "(columns, permutations, nodes; %d timed step(s))"
                  % (args.log_rows, c, C, 1 << args.log_blowup, leaf_rows.bit_length() - 1, args.log_rows + args.log_blowup,
                     leaf_rows.bit_length() - 1, t_lde, t_tree, t_lev, full_lde, full_leaf, full_lev, len(steps)),
        "seconds_per_step_measured": t_lde + t_tree + t_lev, "seconds_full_job": t_full, "sample_cols": c, "sample_leaf_rows": leaf_rows,
        "full_job_seconds": {"extendPol": full_lde, "leaf_hash": full_leaf, "tree_levels": full_lev},
        "poseidon_hash_per_s": (leaf_rows * pc + leaf_rows - 1) / t_tree if t_tree > 0 else None,
        "omp": {"threads": cores, "OMP_PROC_BIND": os.environ.get("OMP_PROC_BIND"), "build": "AVX2" if kind == "reference" else "scalar C"},
    }


def workload_config(args):
    """`config` is the same object in both arms (the driver compares them)."""
    return {"workload": "zkEVM batch-proof stage-1 commit: 2^%d rows x %d cols -> 2^%d rows, extendPol + MerkleTreeGL (BASELINE.json configs[1])"
                        % (args.log_rows, args.cols, args.log_rows + args.log_blowup),
            "l2_policy": "inputs (%.1f GB) far larger than the 126 MB L2; no flush needed" % (8.0 * (1 << args.log_rows) * args.cols / 1e9),
            "seed": hex(SEED),
            "cpu_sample": "the CPU arms time %d of the %d columns at the full row count and extrapolate by component (cpu_baseline.sample)"
                          % (min(CPU_SAMPLE_COLS, args.cols), args.cols)}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    base = run_cpu_baseline(args, budget_s=150.0, nsteps=max(1, args.steps), warmup=max(0, min(args.warmup, 1)))
    keys = ("value", "unit", "cores", "kind", "sample", "extrapolated", "full_job_seconds", "omp")
    line = {
        "impl": "reference", "metric": "LDE+Merkle commit throughput (2^%d x %d, blowup %d)" % (args.log_rows, args.cols, 1 << args.log_blowup),
        "value": base["value"], "unit": "Melem/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": base["seconds_full_job"] * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic", "extrapolated": True,
        "config": workload_config(args),
        "parallelism": "host CPU, OpenMP x%d" % base["cores"],
        "cpu_baseline": {k: base[k] for k in keys},
        "e2e": {"value": base["value"], "unit": "Melem/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "poseidon_hash_per_s": base["poseidon_hash_per_s"], "gpu_launches": 0,
    }
    emit(line)
    return 0


# --------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------
def verify_full_size(ctx, torch, ext_t, tree_t, n, n_ext, ncols, col0, ncols_local, rows0, nrows_local):
    """Spot-check parity at full size against the oracle (outside the timed region):
    two whole columns of the extension, and Merkle paths of sampled leaves up to the root."""
    from oracle import refbind
    from xlayer_prover_b200._capi import synth_matrix
    orc = refbind.best()
    out = {}
    if ext_t is not None and rows0 == 0 and nrows_local == n_ext:
        cols = sorted(set([0, ncols_local - 1]))
        a = np.concatenate([synth_matrix(n, 1, SEED, col0=col0 + c, ncols_total=ncols) for c in cols], axis=1)
        want = orc.extend_pol(a, n_ext)
        got = ext_t.view(n_ext, ncols_local)[:, cols].cpu().numpy().view(np.uint64)
        out["lde_columns_checked"] = [col0 + c for c in cols]
        out["lde_columns_equal"] = bool(np.array_equal(got, want))
    if tree_t is not None and ext_t is not None and ncols_local == ncols:
        tree = tree_t.cpu().numpy().view(np.uint64)
        depth = n_ext.bit_length() - 1
        ok = True
        rng = np.random.default_rng(1)
        for idx in [0, n_ext - 1] + [int(x) for x in rng.integers(0, n_ext, size=14)]:
            row = ext_t.view(n_ext, ncols)[idx].cpu().numpy().view(np.uint64)
            h = orc.linear_hash(row)
            ok &= bool(np.array_equal(h, tree[4 * idx: 4 * idx + 4]))
            off, nn, i = 0, n_ext, idx
            for _ in range(depth):
                sib = tree[off + 4 * (i ^ 1): off + 4 * (i ^ 1) + 4]
                pair = np.concatenate([h, sib]) if i % 2 == 0 else np.concatenate([sib, h])
                h = orc.hash_full_result(np.concatenate([pair, np.zeros(4, dtype=np.uint64)]))[:4]
                off += 4 * nn
                nn //= 2
                i //= 2
            ok &= bool(np.array_equal(h, tree[-4:]))
        out["merkle_paths_checked"] = 16
        out["merkle_paths_equal"] = ok
    out["oracle"] = "reference build" if refbind.reference() else "C port"
    return out


def main_ours(args):
    import torch
    import torch.distributed as dist
    from xlayer_prover_b200 import Context, tree_num_elements
    from xlayer_prover_b200.sharded import ShardedCommit

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        if args.gpus > 1:
            sys.stderr.write("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d\n" % (args.gpus, args.gpus))
            return 2
    if not torch.cuda.is_available():
        sys.stderr.write("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)\n")
        return 2
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n, n_ext, C = 1 << args.log_rows, 1 << (args.log_rows + args.log_blowup), args.cols
    K, W = args.steps, args.warmup
    ctx = Context(local_rank)
    stream = torch.cuda.Stream(device=dev)
    ctx.set_stream(stream.cuda_stream)
    hbm_peak, peak_src = measured_peaks()
    int_peak = ctx.measure_pipe_peaks()

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    result = {}
    with torch.cuda.stream(stream):
        sc = ShardedCommit(ctx, n, n_ext, C, device=dev) if world > 1 else None
        c0, w = (sc.c0, sc.w) if sc else (0, C)
        # ---- inputs resident in HBM (generated on the device, identical for every N) ----
        din = torch.empty(max(1, n * w), dtype=torch.int64, device=dev)
        ctx.synth_fill(din, n, w, SEED, col0=c0, ncols_total=C)
        if sc is None:
            dext = torch.empty(n_ext * C, dtype=torch.int64, device=dev)
            dtree = torch.empty(tree_num_elements(n_ext), dtype=torch.int64, device=dev)

        def step():
            if sc is None:
                return ctx.commit(dext, dtree, din, n, n_ext, C)
            return sc.commit(din)

        for _ in range(W):
            step()
        ctx.get_profile()
        barrier()
        launches0 = ctx.launch_count()
        retries0, redo0 = ctx.ntt_retry_count(), ctx.leaf_redo_rows()
        ctx.set_profiling(True)
        sampler = ClockSampler(local_rank)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        root = None
        for _ in range(K):
            root = step()
        e1.record(stream)
        barrier()
        clocks = sampler.stop()
        ctx.set_profiling(False)
        prof = ctx.get_profile()
        launches = ctx.launch_count() - launches0
        retries, redo = ctx.ntt_retry_count() - retries0, ctx.leaf_redo_rows() - redo0
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms_per_step = float(ms.item()) / K
        result.update(ms_per_step=ms_per_step, launches=launches, prof=prof, clocks=clocks, retries=retries, redo=redo,
                      root=[int(x) for x in (root.cpu().numpy().view(np.uint64) if hasattr(root, "cpu") else root)] if root is not None else None)

        # ---- parity spot-check at full size (untimed) ----
        verify = None
        if not args.no_verify and rank == 0:
            try:
                if sc is None:
                    verify = verify_full_size(ctx, torch, dext, dtree, n, n_ext, C, 0, C, 0, n_ext)
                else:
                    verify = {"note": "N > 1: the tree is assembled on rank 0; slabs and nodes are compared with the N = 1 result by tests/test_gpu_parity.py::test_sharded_commit_two_gpus"}
                known = KNOWN_ROOTS.get((args.log_rows, C, args.log_blowup))
                verify["root_equals_n1"] = (result["root"] == known) if (known and result["root"]) else None
                verify["root_n1_source"] = "committed root of the N = 1 run of this workload (bench.py KNOWN_ROOTS), itself checked against the reference build at N = 1" if known else "no committed root for this shape"
            except Exception as ex:                                     # never let the checker hide the measurement
                verify = {"error": repr(ex)}
        result["verify"] = verify

        # ---- e2e: host buffers through the C ABI ----
        e2e = None
        if not args.no_e2e:
            if sc is None:
                del dext, dtree, din
            torch.cuda.empty_cache()
            try:
                e2e = run_e2e(args, ctx, torch, dist, dev, stream, sc, rank, world, n, n_ext, C, c0, w, barrier)
            except Exception as ex:                                     # keep the device-resident measurement
                e2e = {"value": None, "unit": "Melem/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0, "note": "failed: %r" % (ex,)}
        if e2e and e2e.get("root") is not None and rank == 0:
            e2e["root_equals_device_path"] = e2e.pop("root") == result["root"]
        elif e2e:
            e2e.pop("root", None)
        result["e2e_resident"] = e2e.pop("e2e_resident", None) if e2e else None
        result["e2e"] = e2e

    if rank == 0:
        perms = perms_per_commit(n_ext, C)
        sec = result["ms_per_step"] / 1e3
        prof = result["prof"]
        leaf_ms = prof["leaf_hash"][0] / max(1, prof["leaf_hash"][1])
        R = n_ext // world
        leaf_bytes = 8.0 * C * R + 32.0 * R                              # algorithmic bytes of one leaf-hash launch
        leaf_gbs = leaf_bytes / (leaf_ms * 1e-3) / 1e9 if leaf_ms > 0 else None
        leaf_perms = R * ((C + 7) // 8)
        leaf_pps = leaf_perms / (leaf_ms * 1e-3) if leaf_ms > 0 else None
        clk_hz = (result["clocks"].get("sm_mhz") or 1965.0) * 1e6
        alg_peak = 1.0 / (POS_ALG_IMAD_WIDE / int_peak["imad_wide_lane_ops_per_s"] + POS_ALG_DFMA / int_peak["dfma_lane_ops_per_s"])
        lde_ms = (prof["intt"][0] + prof["ntt_ext"][0]) / K
        wcols = w
        lde_bytes = 8.0 * wcols * (n + n_ext)
        line = {
            "metric": "LDE+Merkle commit throughput (2^%d x %d, blowup %d)" % (args.log_rows, C, 1 << args.log_blowup),
            "value": n * C / sec / 1e6, "unit": "Melem/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": result["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": workload_config(args),
            "parallelism": "single GPU" if world == 1 else "column-sharded LDE x%d whose last NTT pass stores into the peers' row slabs over NVLink (fused exchange) -> row-sharded Merkle, top %d levels on rank 0" % (world, world.bit_length() - 1),
            "commit_ms": result["ms_per_step"], "poseidon_hash_per_s": perms / sec,
            "leaf_hash_rows_per_s": (R / (leaf_ms * 1e-3)) if leaf_ms > 0 else None,
            "stage_ms": {"intt": prof["intt"][0] / K, "ntt_ext": prof["ntt_ext"][0] / K, "leaf_hash": prof["leaf_hash"][0] / K,
                         "tree_levels": prof["tree_levels"][0] / K, "note": "rank 0, per step, CUDA events on the launching stream"},
            "ntt_retries": result["retries"], "leaf_redo_rows": result["redo"],
            "roofline": {"kernel": "k_linear_hash_rows", "bound": "int", "achieved": leaf_pps, "peak": alg_peak,
                         "unit": "Poseidon permutations/s", "frac": (leaf_pps / alg_peak) if (leaf_pps and alg_peak) else None,
                         "traffic": LEAF_DRAM_BYTES_PER_ROW_660 * R if C == 660 else None,
                         "traffic_note": "ncu dram bytes per launch, scaled from the 2^19-row capture by rows (3.0 % over the algorithmic bytes 8*C + 32 per row)" if C == 660 else None,
                         "launch_ms": leaf_ms, "permutations_per_launch": leaf_perms,
                         "peak_source": "algorithmic integer ceiling: %d IMAD.WIDE + %d DFMA lane-instructions per permutation at the rates "
                                        "glp_measure_pipe_peaks measured in this run (%.3g and %.3g lane-ops/s)"
                                        % (POS_ALG_IMAD_WIDE, POS_ALG_DFMA, int_peak["imad_wide_lane_ops_per_s"], int_peak["dfma_lane_ops_per_s"]),
                         "note": "dominant kernel by time (%.0f %% of the step); integer work, not HBM: see roofline_hbm for the byte figure and "
                                 "roofline_sass / roofline_issue for how close the kernel runs to its own instruction stream"
                                 % (100.0 * (prof["leaf_hash"][0] / K) / result["ms_per_step"])},
            "roofline_hbm": {"kernel": "k_linear_hash_rows", "bound": "hbm", "achieved": leaf_gbs, "peak": hbm_peak, "unit": "GB/s",
                             "frac": (leaf_gbs / hbm_peak) if leaf_gbs else None, "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": leaf_bytes},
            "roofline_sass": {"kernel": "k_linear_hash_rows", "bound": "multiplier pipes of the kernel's own SASS (fmaheavy IMAD + FP64)",
                              "achieved": leaf_pps, "peak": 148 * 4 * 32 * clk_hz / POS_BOUND_CYCLES, "unit": "Poseidon permutations/s",
                              "frac": leaf_pps / (148 * 4 * 32 * clk_hz / POS_BOUND_CYCLES) if leaf_pps else None,
                              "pipe_cycles_per_warp_permutation": POS_PIPE_CYCLES, "sass_instructions_per_permutation": POS_SASS,
                              "source": "tools/sass_leaf.py on the shipped binary (profiles/r02_sass_mix_leaf.txt)"},
            "roofline_issue": {"kernel": "k_linear_hash_rows", "bound": "mixed integer issue rate (IMAD + IADD3 interleaved, measured in this run)",
                               "achieved": leaf_pps * POS_SASS["instructions"] if leaf_pps else None,
                               "peak": int_peak["interleaved_lane_ops_per_s"], "unit": "lane-instructions/s",
                               "frac": leaf_pps * POS_SASS["instructions"] / int_peak["interleaved_lane_ops_per_s"] if leaf_pps else None,
                               # IMAD.WIDE is a half-rate instruction (measured alone at half the IMAD rate): counted as two issue slots
                               "frac_imad_wide_as_two_slots": leaf_pps * (POS_SASS["instructions"] + POS_SASS["imad_wide"]) / int_peak["interleaved_lane_ops_per_s"]
                               if leaf_pps else None,
                               "note": "the B200 issues ~70 integer lane-instructions per clock per SM whatever the IMAD/ALU mix (IMAD alone 60, IADD3 alone 61, "
                                       "interleaved 70): the kernel's stream runs at that rate, so only a shorter stream makes it faster"},
            "measured_lane_ops_per_s": int_peak,
            "roofline_lde": {"kernels": "k_ntt_pass: 3 INTT + 3 NTT tile passes per column block", "bound": "hbm", "achieved": lde_bytes / (lde_ms * 1e-3) / 1e9 if lde_ms > 0 else None,
                             "peak": hbm_peak, "unit": "GB/s", "frac": (lde_bytes / (lde_ms * 1e-3) / 1e9 / hbm_peak) if lde_ms > 0 else None,
                             "algorithmic_bytes": lde_bytes, "design_traffic_bytes": 8.0 * wcols * (5 * n + 3 * n_ext), "lde_ms": lde_ms},
            "clocks": result["clocks"], "gpu_launches": result["launches"], "e2e": result["e2e"], "e2e_resident": result.get("e2e_resident"),
            "root": ["%#018x" % x for x in result["root"]] if result["root"] else None, "verify": result["verify"],
        }
        if world == 1 and not args.no_cpu_baseline:
            try:
                base = run_cpu_baseline(args, args.cpu_budget_s)
                line["cpu_baseline"] = {k: base[k] for k in ("value", "unit", "cores", "kind", "sample", "extrapolated", "full_job_seconds", "omp")}
            except Exception as ex:
                line["cpu_baseline"] = {"error": repr(ex)}
        emit(line)
    if sc is not None:
        sc.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()
    return 0


def run_e2e(args, ctx, torch, dist, dev, stream, sc, rank, world, n, n_ext, C, c0, w, barrier):
    """Same commit, host buffers: per step H2D of this rank's trace columns and D2H of its extended rows +
    (rank 0) every Merkle node, all inside the timed region."""
    from xlayer_prover_b200 import tree_num_elements
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 64 << 30
    K, W = args.steps, args.warmup
    R = n_ext // world
    in_bytes, ext_bytes, tree_bytes = 8 * n * w, 8 * R * C, 8 * tree_num_elements(n_ext)
    need = (in_bytes + ext_bytes) * (world if world > 1 else 1) + tree_bytes
    full = need < 0.8 * avail
    note = "H2D trace, D2H extended matrix + all Merkle nodes + root"
    if not full:
        need = in_bytes * (world if world > 1 else 1) + tree_bytes
        if need > 0.8 * avail:
            return {"value": None, "unit": "Melem/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                    "note": "skipped: host has %.0f GB available, needs %.0f GB of pinned memory" % (avail / 1e9, need / 1e9)}
        note = "host RAM too small to pin the %.0f GB extended matrix: H2D trace, D2H Merkle nodes + root only" % (ext_bytes * world / 1e9)
    # pinned host buffers straight from cudaHostAlloc (torch's pinned allocator rounds sizes up to a power of two)
    import ctypes

    def pinned(nwords):
        p = ctx.host_alloc(8 * max(1, nwords))
        t = torch.frombuffer((ctypes.c_char * (8 * max(1, nwords))).from_address(p), dtype=torch.int64)
        return p, t
    hin_p, hin = pinned(n * w)
    hext_p, hext = pinned(R * C) if full else (None, None)
    htree_p, htree = pinned(tree_num_elements(n_ext)) if rank == 0 else (None, None)
    # fill the pinned input (device generator -> host), then drop the device copy
    tmp = torch.empty(max(1, n * w), dtype=torch.int64, device=dev)
    ctx.synth_fill(tmp, n, w, SEED, col0=c0, ncols_total=C)
    ctx.memcpy(hin_p, tmp, 8 * n * w)
    del tmp
    torch.cuda.empty_cache()

    if sc is None:
        def step():
            return ctx.commit(hext_p, htree_p, hin_p, n, n_ext, C)
    else:
        def step():
            return sc.commit_host(hin_p, hext, htree if rank == 0 else None)
    for _ in range(min(W, 3)):
        step()
    barrier()
    t0 = time.perf_counter()
    root = None
    for _ in range(K):
        root = step()
    barrier()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    sec = float(dt.item()) / K
    resident = None
    if sc is None and not args.no_e2e_resident:
        # the same commit with the RESULTS kept on the device: per step H2D of the trace, then only what a prover
        # reads back -- the root and the 128 query proofs (row + Merkle path) of this tree, gathered in one launch
        # (glp_merkle_group_proofs, friProve.cpp:150-170).  Reported next to, not instead of, the host-buffer figure.
        try:
            if hext_p:
                ctx.host_free(hext_p)
                hext_p = None
            ctx.release_scratch()                                   # the host-pointer leg's staging buffers (~100 GB) would not fit next to dext
            torch.cuda.empty_cache()
            dext = torch.empty(n_ext * C, dtype=torch.int64, device=dev)
            dtree = torch.empty(tree_num_elements(n_ext), dtype=torch.int64, device=dev)
            nq = 128
            idx = np.random.default_rng(3).integers(0, n_ext, size=nq, dtype=np.uint64)
            depth = n_ext.bit_length() - 1
            proofs = np.zeros((nq, C + 4 * depth), dtype=np.uint64)

            def rstep():
                r = ctx.commit(dext, dtree, hin_p, n, n_ext, C)
                ctx.merkle_group_proofs(proofs, dext, dtree, C, n_ext, idx)
                return r
            for _ in range(2):
                rstep()
            barrier()
            t0r = time.perf_counter()
            for _ in range(K):
                rroot = rstep()
            barrier()
            rsec = (time.perf_counter() - t0r) / K
            resident = {"value": n * C / rsec / 1e6, "unit": "Melem/s", "ms_per_step": rsec * 1e3, "h2d_bytes_per_step": 8 * n * C,
                        "d2h_bytes_per_step": 32 + proofs.nbytes, "queries": nq,
                        "root_equals_host_path": [int(x) for x in np.asarray(rroot).view(np.uint64)] == [int(x) for x in np.asarray(root).view(np.uint64)],
                        "note": "H2D trace; extension and tree stay in HBM; D2H root + 128 query proofs (one glp_merkle_group_proofs call)"}
            del dext, dtree
            torch.cuda.empty_cache()
        except Exception as ex:
            resident = {"value": None, "note": "failed: %r" % (ex,)}
    for p in (hin_p, hext_p, htree_p):
        if p:
            ctx.host_free(p)
    return {"e2e_resident": resident, "value": n * C / sec / 1e6, "unit": "Melem/s", "ms_per_step": sec * 1e3,
            "h2d_bytes_per_step": 8 * n * C, "d2h_bytes_per_step": (8 * n_ext * C if full else 0) + tree_bytes + 32,
            "note": note, "pinned": True,
            "root": [int(x) for x in np.asarray(root).view(np.uint64)] if (root is not None and not hasattr(root, "cpu")) else
                    ([int(x) for x in root.cpu().numpy().view(np.uint64)] if root is not None else None)}


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries write banners to stdout (NCCL prints its version there); the contract is ONE JSON line on stdout.
    Everything else goes to stderr until emit() restores the descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    if _REAL_STDOUT is not None:
        os.dup2(_REAL_STDOUT, 1)
    print(json.dumps(line), flush=True)
    if _REAL_STDOUT is not None:
        os.dup2(2, 1)                              # anything printed during teardown goes to stderr again


def main():
    args = parse_args()
    host_cores()
    if args.impl == "reference":
        # before libgomp initialises: every core the process may use, threads spread over them.
        # torch.distributed.run exports OMP_NUM_THREADS=1 for its workers; rank 0 of the reference arm is the only worker.
        os.environ["OMP_NUM_THREADS"] = str(host_cores())
        os.environ.setdefault("OMP_PROC_BIND", "spread")
    quiet_stdout()
    if args.impl == "reference":
        return main_reference(args)
    return main_ours(args)


if __name__ == "__main__":
    sys.exit(main())
