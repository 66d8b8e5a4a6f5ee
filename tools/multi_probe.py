#!/usr/bin/env python3
"""glp_multi_commit on two devices at a moderate size (2^20 x 64 -> 2^21), for ncu: the last NTT pass of each device stores half of
its rows into the other device's slab.  usage: multi_probe.py [log_rows] [cols]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from xlayer_prover_b200 import MultiContext, synth_matrix, tree_num_elements  # noqa: E402

log_rows = int(sys.argv[1]) if len(sys.argv) > 1 else 20
cols = int(sys.argv[2]) if len(sys.argv) > 2 else 64
n, n_ext = 1 << log_rows, 2 << log_rows
a = synth_matrix(n, cols, 7)
m = MultiContext([0, 1])
ext = np.zeros((n_ext, cols), dtype=np.uint64)
tree = np.zeros(tree_num_elements(n_ext), dtype=np.uint64)
for _ in range(2):
    root = m.commit(ext, tree, a, n, n_ext, cols)
print("root[0] = %#x" % int(root[0]))
m.close()
