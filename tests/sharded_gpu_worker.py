"""One rank of the multi-GPU commit on REAL devices (NCCL, CUDA IPC, the fused NVLink scatter), launched by
tests/test_gpu_parity.py::test_sharded_commit_two_gpus under torch.distributed.run.  Every rank checks its own row slab of the
extension against the single-GPU result of the same library (computed by rank 0 on its own device and broadcast as a file)
and rank 0 checks every Merkle node and the root against it and against the oracle."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)


def main():
    out_dir, log_n, blow, ncols = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    from xlayer_prover_b200 import Context, tree_num_elements
    from xlayer_prover_b200.sharded import ShardedCommit
    ctx = Context(local_rank)
    n, n_ext = 1 << log_n, 1 << (log_n + blow)
    seed = 0xABCDEF
    # N = 1 result on rank 0's device (the product path, device-resident), written for every rank to read
    ref_path = os.path.join(out_dir, "n1.npz")
    if rank == 0:
        din = torch.empty(n * ncols, dtype=torch.int64, device=dev)
        ctx.synth_fill(din, n, ncols, seed)
        dext = torch.empty(n_ext * ncols, dtype=torch.int64, device=dev)
        dtree = torch.empty(tree_num_elements(n_ext), dtype=torch.int64, device=dev)
        ctx.commit(dext, dtree, din, n, n_ext, ncols)
        np.savez(ref_path, ext=dext.cpu().numpy().view(np.uint64).reshape(n_ext, ncols), tree=dtree.cpu().numpy().view(np.uint64))
        del din, dext, dtree
    dist.barrier()
    ref = np.load(ref_path)
    stream = torch.cuda.Stream(device=dev)
    ctx.set_stream(stream.cuda_stream)
    with torch.cuda.stream(stream):
        sc = ShardedCommit(ctx, n, n_ext, ncols, device=dev)
        assert sc.fused, "the fused IPC scatter is the path under test"
        din = torch.empty(max(1, n * sc.w), dtype=torch.int64, device=dev)
        ctx.synth_fill(din, n, sc.w, seed, col0=sc.c0, ncols_total=ncols)
        root = sc.commit(din)
        torch.cuda.synchronize(dev)
        slab = sc.slab.cpu().numpy().view(np.uint64).reshape(sc.R, ncols)
        ok_slab = bool(np.array_equal(slab, ref["ext"][rank * sc.R:(rank + 1) * sc.R]))
        ok_tree = True
        if rank == 0:
            tree = sc.tree.cpu().numpy().view(np.uint64)
            ok_tree = bool(np.array_equal(tree, ref["tree"])) and bool(np.array_equal(root.cpu().numpy().view(np.uint64), ref["tree"][-4:]))
        flags = torch.tensor([int(ok_slab), int(ok_tree)], dtype=torch.int32, device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        sc.close()
    if rank == 0:
        with open(os.path.join(out_dir, "result.txt"), "w") as f:
            f.write("slabs_equal=%d tree_equal=%d world=%d\n" % (int(flags[0]), int(flags[1]), world))
    dist.barrier()
    dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
