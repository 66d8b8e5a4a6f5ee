"""Multi-GPU commit (xlayer_prover_b200/sharded.py) on the CPU: world_size 1, 2, 4 and 8 gloo ranks, each driving the
host-thread emulator build of the kernels.  The extension slabs, every Merkle node and the root must equal the
single-device oracle result for every world size (results identical for every G)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from cases import check_equal

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def emu_built():
    subprocess.run([os.path.join(HERE, "cuemu", "build_emu.sh")], check=True, capture_output=True)


def test_column_bounds_and_assemble():
    from xlayer_prover_b200.sharded import assemble_tree, column_bounds
    from xlayer_prover_b200._capi import tree_num_elements
    for ncols in (1, 7, 8, 20, 660, 665, 751):
        for g in (1, 2, 4, 8):
            b = column_bounds(ncols, g)
            assert b[0] == 0 and b[-1] == ncols and all(x <= y for x, y in zip(b, b[1:]))
            assert all(x % 8 == 0 for x in b[:-1])           # every 8-wide sponge block has one owner
    # level-major re-layout: node k of level l of the full tree = node (k mod n_l/G) of level l of subtree k // (n_l/G)
    n_ext, g = 16, 4
    subs = [torch.arange(tree_num_elements(n_ext // g), dtype=torch.int64) + 1000 * r for r in range(g)]
    full, off = assemble_tree(subs, n_ext, g)
    assert off == 4 * (16 + 8)
    assert full[0:16].tolist() == list(range(16)) and full[16:32].tolist() == [1000 + i for i in range(16)]
    assert full[64:72].tolist() == list(range(16, 24))      # level 1: each subtree contributes 2 nodes
    assert full[off: off + 16].tolist() == [24, 25, 26, 27, 1024, 1025, 1026, 1027, 2024, 2025, 2026, 2027, 3024, 3025, 3026, 3027]


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_sharded_commit_gloo(emu_built, oracle, tmp_path, world):
    from xlayer_prover_b200._capi import synth_matrix
    log_n, blow, ncols = 5, 1, 20
    out = str(tmp_path / "r0.npz")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + world), os.path.join(HERE, "sharded_worker.py"), out, str(log_n), str(blow), str(ncols)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    got = np.load(out)
    a = synth_matrix(1 << log_n, ncols, 0xABCDEF)
    want_ext = oracle.extend_pol(a, 1 << (log_n + blow))
    want_tree = oracle.merkletree(want_ext)
    check_equal(got["ext"], want_ext, "row slabs, world %d" % world)
    check_equal(got["tree"], want_tree, "assembled tree, world %d" % world)
    check_equal(got["root"], want_tree[-4:], "root, world %d" % world)
