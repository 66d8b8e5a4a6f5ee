"""One rank of the multi-GPU commit on the CPU: gloo for the collectives, the host-thread emulator build of the
kernel sources (tests/cuemu) as the 'device'.  Launched by tests/test_sharded.py; writes rank 0's result to a file."""
import ctypes
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)


def main():
    out_path, log_n, blow, ncols = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    from xlayer_prover_b200._capi import Context, GlpLib, synth_matrix
    from xlayer_prover_b200.sharded import ShardedCommit, to_numpy_u64
    ctx = Context(0, GlpLib(os.path.join(HERE, "cuemu", "libglprover_emu.so")))
    keep = []

    def alloc(nwords):                       # emulator "device" memory viewed as a CPU tensor
        p = ctx.device_alloc(8 * max(1, nwords))
        keep.append(p)
        return torch.frombuffer((ctypes.c_char * (8 * max(1, nwords))).from_address(p), dtype=torch.int64)

    n, n_ext = 1 << log_n, 1 << (log_n + blow)
    sc = ShardedCommit(ctx, n, n_ext, ncols, device="cpu", alloc=alloc)
    a = synth_matrix(n, sc.w, 0xABCDEF, col0=sc.c0, ncols_total=ncols)           # this rank's column slice only
    din = alloc(n * sc.w)
    din[: n * sc.w] = torch.from_numpy(a.view(np.int64).reshape(-1).copy())
    root = sc.commit(din)
    slab = to_numpy_u64(sc.slab).reshape(sc.R, ncols).copy()
    tree = to_numpy_u64(sc.tree).copy() if rank == 0 else None
    # the host-pointer flavour (what bench.py's e2e leg runs at N > 1) must give the same slab, tree and root
    hext = torch.zeros(sc.R * ncols, dtype=torch.int64)
    htree = torch.zeros(sc.subtree.numel() * 0 + (2 * n_ext - 1) * 4, dtype=torch.int64) if rank == 0 else None
    root2 = sc.commit_host(np.ascontiguousarray(a), hext, htree)
    assert np.array_equal(to_numpy_u64(hext).reshape(sc.R, ncols), slab), "commit_host slab differs"
    if rank == 0:
        assert np.array_equal(to_numpy_u64(root2), to_numpy_u64(root)) and np.array_equal(to_numpy_u64(htree), tree), "commit_host tree differs"
    slabs = [None] * world
    dist.gather_object(slab, slabs if rank == 0 else None, dst=0)
    if rank == 0:
        np.savez(out_path, root=to_numpy_u64(root), tree=to_numpy_u64(sc.tree), ext=np.concatenate(slabs, axis=0))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
