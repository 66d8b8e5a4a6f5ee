"""Randomised shapes through the C ABI against the oracle (hypothesis, derandomised so every run draws the same cases):
ragged column counts, every blow-up the reference uses, host and device pointers, values over the whole u64 range
(non-canonical words included).  On the CPU the same draws run on the host-thread emulator build of the kernels."""
import os
import subprocess

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from cases import check_commit, check_equal, edge_matrix, run_extend, run_merkletree, run_ntt

HERE = os.path.dirname(os.path.abspath(__file__))
COMMON = dict(deadline=None, derandomize=True, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])

shape = st.tuples(st.integers(0, 9), st.integers(1, 41), st.integers(0, 4), st.integers(0, 2 ** 32 - 1), st.booleans())


def one_case(ctx, oracle, logn, ncols, logb, seed, device):
    rng = np.random.default_rng(seed)
    n = 1 << logn
    if (n << logb) == 4 and n < 4:
        # the reference build itself aborts inside extendPol for N = 1 or 2 extended to exactly 4 rows (its phase
        # split degenerates); these two shapes are checked against the pinned C restatement instead
        from oracle import refbind
        oracle = refbind.port()
    a = edge_matrix(rng, n, ncols)
    check_commit(ctx, oracle, a, n << logb, device=device)
    check_equal(run_ntt(ctx, a, True, False), oracle.canonical(oracle.ntt(a, inverse=True)), "INTT %s" % ((n, ncols),))
    check_equal(run_merkletree(ctx, a), oracle.canonical(oracle.merkletree(a)), "merkletree %s" % ((n, ncols),))


@pytest.fixture(scope="module")
def ectx():
    subprocess.run([os.path.join(HERE, "cuemu", "build_emu.sh")], check=True, capture_output=True)
    from xlayer_prover_b200._capi import Context, GlpLib
    os.environ["GLP_NTT_MAX_LOGQ"] = "4"          # 2^9 rows = three tile passes on the emulator
    try:
        c = Context(0, GlpLib(os.path.join(HERE, "cuemu", "libglprover_emu.so")))
    finally:
        os.environ.pop("GLP_NTT_MAX_LOGQ", None)
    yield c
    c.close()


@settings(max_examples=20, **COMMON)
@given(shape)
def test_random_shapes_emulator(ectx, oracle, s):
    logn, ncols, logb, seed, device = s
    one_case(ectx, oracle, min(logn, 7), ncols, min(logb, 3), seed, device)


@pytest.mark.gpu
@settings(max_examples=int(os.environ.get("GLP_FUZZ_EXAMPLES", "60")), **COMMON)      # GLP_FUZZ_EXAMPLES=1000 for a long run
@given(st.tuples(st.integers(0, 15), st.integers(1, 130), st.integers(0, 4), st.integers(0, 2 ** 32 - 1), st.booleans()))
def test_random_shapes_gpu(ctx, oracle, s):
    logn, ncols, logb, seed, device = s
    if (1 << (logn + logb)) * ncols > 1 << 24:   # keep the oracle side in seconds
        ncols = max(1, (1 << 24) >> (logn + logb))
    one_case(ctx, oracle, logn, ncols, logb, seed, device)
