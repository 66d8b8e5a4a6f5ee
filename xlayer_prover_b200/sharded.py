"""Multi-GPU commit: column-sharded LDE -> all-to-all over NVLink -> row-sharded Merkle subtrees -> top levels on rank 0.

One process per GPU (torch.distributed; NCCL on the B200 box, gloo in the CPU tests).  The path shards
exactly as BASELINE.json's north_star prescribes:

  1. columns are independent polynomials: rank g extends its own column slice [c_g, c_{g+1}) with
     glp_extend_pol_pitched (boundaries are multiples of 8, so every 8-wide sponge block has one owner);
  2. ONE exchange step: rank g sends rows [j*R, (j+1)*R) (R = N_ext/G) of its slice to rank j
     (all_to_all_single), and each rank interleaves what it received into a reference-layout
     [R x C] row slab of the extended matrix;
  3. rank j hashes its R leaf rows and builds its subtree (glp_merkletree on the slab);
     With fused=True (the default on CUDA) steps 1 and 2 are ONE pass over the data: every rank maps its peers' row
     slabs (CUDA IPC over NVLink) and the LDE's last NTT pass stores each tile straight into the owner's slab at this
     rank's column offset (glp_extend_pol_scatter) -- no all-to-all, no receive buffer, no interleave pass; the only
     collective left is a one-word all-reduce that orders the peers' stores before the hashing;
  4. every rank copies the levels of its subtree straight into their slices of the reference's level-major `nodes` layout
     (poseidon_goldilocks.cpp:283-300) in rank 0's tree, mapped over CUDA IPC (glp_place_subtree: ~25 peer copies, no gather,
     no re-layout pass), and rank 0 hashes the top log2(G) levels (glp_merkle_levels).  The unfused path (gloo tests) gathers
     the subtrees and re-lays them on rank 0.

torch is plumbing only: it owns the buffers, the NCCL communicator and the current stream; all arithmetic
is in libglprover.so.  Results are identical for every G (tests/test_sharded.py).
"""
import numpy as np
import torch
import torch.distributed as dist

from ._capi import tree_num_elements


def column_bounds(ncols, world):
    """c_0..c_G with every boundary a multiple of 8 (except the last, = ncols)."""
    nblk = (ncols + 7) // 8
    return [min(ncols, 8 * ((g * nblk) // world)) for g in range(world)] + [ncols]


def assemble_tree(subtrees, n_ext, world, empty=None):
    """Re-lay G row-sharded subtrees (each (2R-1)*4 words, level-major) into the first
    levels of the full level-major tree; returns (full_tree, offset_of_level_with_G_nodes)."""
    R = n_ext // world
    full = (empty(tree_num_elements(n_ext)) if empty is not None else
            torch.empty(tree_num_elements(n_ext), dtype=torch.int64, device=subtrees[0].device))
    off_full, off_sub, n_full, n_sub = 0, 0, n_ext, R
    while n_sub >= 1:
        for g, t in enumerate(subtrees):
            full[off_full + 4 * g * n_sub: off_full + 4 * (g + 1) * n_sub] = t[off_sub: off_sub + 4 * n_sub]
        if n_sub == 1:
            break
        off_full += 4 * n_full
        off_sub += 4 * n_sub
        n_full //= 2
        n_sub //= 2
    return full, off_full


class ShardedCommit:
    """commit(in_slice) for one rank.  `ctx` is this rank's glprover Context; `device` the torch device the
    buffers live on ('cuda:N', or 'cpu' with the emulator backend in the gloo tests)."""

    def __init__(self, ctx, n, n_ext, ncols, group=None, device=None, alloc=None, fused=None):
        self.ctx, self.n, self.n_ext, self.ncols = ctx, n, n_ext, ncols
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        assert n_ext % self.world == 0 and (self.world & (self.world - 1)) == 0, "world size must be a power of two dividing n_ext"
        self.device = torch.device(device if device is not None else "cuda:%d" % ctx.device)
        self.bounds = column_bounds(ncols, self.world)
        self.c0, self.c1 = self.bounds[self.rank], self.bounds[self.rank + 1]
        self.w = self.c1 - self.c0
        self.R = n_ext // self.world
        i64 = dict(dtype=torch.int64, device=self.device)
        # alloc(nwords) -> int64 tensor; the gloo tests pass one that carves "device" memory out of the emulator
        empty = alloc if alloc is not None else (lambda nwords: torch.empty(nwords, **i64))
        self.fused = (self.device.type == "cuda" and self.world > 1) if fused is None else bool(fused)
        if self.fused:
            self._init_fused()
            self.subtree = empty(tree_num_elements(self.R))
            self._empty = empty
            self.tree = None
            return
        # buffer A holds my columns for all rows (LDE output), then -- after the exchange -- my rows for all
        # columns (the slab the Merkle stage reads); buffer B receives the all-to-all.  Peak = A + B.
        self.buf_a = empty(max(1, n_ext * self.w, self.R * ncols))
        self.ext_cols = self.buf_a[: max(1, n_ext * self.w)]
        self.slab = self.buf_a[: self.R * ncols]
        self.recv = empty(self.R * ncols) if self.world > 1 else None
        self.subtree = empty(tree_num_elements(self.R))
        self._empty = empty
        self.tree = None

    def _init_fused(self):
        """My row slab comes straight from cudaMalloc (so it can be exported), every peer's slab is mapped here."""
        R, C, G = self.R, self.ncols, self.world
        self.slab_ptr = self.ctx.device_alloc(8 * max(1, R * C))
        self.slab = _tensor_from_ptr(self.slab_ptr, max(1, R * C), self.device)[: R * C]
        handles = [None] * G
        dist.all_gather_object(handles, self.ctx.ipc_export(self.slab_ptr), group=self.group)
        self.parts = [self.slab_ptr if g == self.rank else self.ctx.ipc_import(handles[g]) for g in range(G)]
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.ext_cols = self.recv = None
        # the full tree lives on rank 0 (cudaMalloc, exported); every rank maps it and writes its subtree's levels straight
        # into their slices of the reference's level-major layout (glp_place_subtree) -- no gather, no re-layout pass
        nt = tree_num_elements(self.n_ext)
        handle = [None]
        if self.rank == 0:
            self.tree_ptr = self.ctx.device_alloc(8 * nt)
            handle[0] = self.ctx.ipc_export(self.tree_ptr)
        dist.broadcast_object_list(handle, src=0, group=self.group)
        if self.rank != 0:
            self.tree_ptr = self.ctx.ipc_import(handle[0])
        self._tree_view = _tensor_from_ptr(self.tree_ptr, nt, self.device) if self.rank == 0 else None
        self._entry_fence = True

    def close(self):
        if getattr(self, "fused", False) and getattr(self, "parts", None):
            if self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
            if dist.is_initialized():
                dist.barrier(group=self.group)           # nobody unmaps while a peer may still be storing
            for g, p in enumerate(self.parts):
                if g != self.rank:
                    self.ctx.ipc_close(p)
            if self.rank != 0:
                self.ctx.ipc_close(self.tree_ptr)
            if dist.is_initialized():
                dist.barrier(group=self.group)           # rank 0 frees the tree after everyone has unmapped it
            self.slab = self.tree = self._tree_view = None
            self.ctx.device_free(self.slab_ptr)
            if self.rank == 0:
                self.ctx.device_free(self.tree_ptr)
            self.parts = None

    def _fence(self):
        """Stream-ordered rendezvous of all ranks (a one-word all-reduce): peers' stores before it are complete and visible after it."""
        dist.all_reduce(self._flag, group=self.group)

    def bind_stream(self):
        """Launch the library's kernels on torch's current stream so they order with the collectives."""
        if self.device.type == "cuda":
            self.ctx.set_stream(torch.cuda.current_stream(self.device).cuda_stream)

    def extend(self, in_slice, in_pitch=None):
        """Step 1: LDE of my column slice.  in_slice: [n x w] (pitch in_pitch, default dense)."""
        if self.fused:
            if self._entry_fence:               # every rank has finished reading its slab; after a complete commit the tree fence
                self._fence()                   # of finish() has already said so
            self._entry_fence = True
            if self.w:
                self.ctx.extend_pol_scatter(self.parts, self.ncols, self.c0, in_slice, in_pitch or self.w, self.n_ext, self.n, self.w)
            return
        if self.w:
            self.ctx.extend_pol_pitched(self.ext_cols, self.w, in_slice, in_pitch or self.w, self.n_ext, self.n, self.w)

    def exchange(self):
        """Step 2: the one collective, then interleave into the row slab."""
        R, C, G = self.R, self.ncols, self.world
        if self.fused:
            self._fence()                       # all ranks' last-pass stores into my slab have landed
            return
        if G == 1:
            return                              # slab and ext_cols are the same memory
        in_splits = [R * self.w] * G
        out_splits = [R * (self.bounds[s + 1] - self.bounds[s]) for s in range(G)]
        dist.all_to_all_single(self.recv, self.ext_cols[: self.n_ext * self.w], out_splits, in_splits, group=self.group)
        slab2d = self.slab.view(R, C)
        off = 0
        for s in range(G):
            ws = self.bounds[s + 1] - self.bounds[s]
            if ws:
                slab2d[:, self.bounds[s]: self.bounds[s + 1]].copy_(self.recv[off: off + R * ws].view(R, ws))
            off += R * ws

    def merkle(self):
        """Step 3: my subtree."""
        self.ctx.merkletree(self.subtree, self.slab, self.ncols, self.R, 1)

    def finish(self):
        """Step 4: gather subtrees on rank 0, assemble, hash the top log2(G) levels.  Returns root on rank 0."""
        G = self.world
        if G == 1:
            self.tree = self.subtree
            return self.tree[-4:].clone()
        if self.fused:
            self.ctx.place_subtree(self.tree_ptr, self.subtree, self.R, G, self.rank)
            self._fence()                       # every subtree's levels are in rank 0's tree; every rank is done with its slab
            self._entry_fence = False
            if self.rank != 0:
                return None
            self.tree = self._tree_view
            off_top, nn = 0, self.n_ext
            while nn > G:
                off_top += 4 * nn
                nn //= 2
            self.ctx.merkle_levels(self.tree[off_top:], G)
            return self.tree[-4:].clone()
        gathered = [torch.empty_like(self.subtree) for _ in range(G)] if self.rank == 0 else None
        dist.gather(self.subtree, gathered, dst=0, group=self.group)
        if self.rank != 0:
            return None
        self.tree, off_top = assemble_tree(gathered, self.n_ext, G, self._empty)
        self.ctx.merkle_levels(self.tree[off_top:], G)
        return self.tree[-4:].clone()

    def commit(self, in_slice, in_pitch=None):
        self.bind_stream()
        self.extend(in_slice, in_pitch)
        self.exchange()
        self.merkle()
        return self.finish()


    def commit_host(self, host_in, host_ext=None, host_tree=None):
        """The same commit with this rank's column slice in (pinned) HOST memory and the results wanted on the host:
        host_in = address or array of the [n x w] slice, host_ext = pinned tensor for this rank's [R x C] row slab,
        host_tree = pinned tensor for the whole tree (rank 0).  Uploads overlap the LDE column block by column block
        (glp_extend_pol's streamed path); the slab's download runs on a side stream while the leaves are hashed."""
        self.bind_stream()
        if self.fused:
            self.extend(host_in)                # glp_extend_pol_scatter streams a host input block by block
        elif self.w:
            self.ctx.extend_pol(self.ext_cols, host_in, self.n_ext, self.n, self.w)
        self.exchange()
        cuda = self.device.type == "cuda"
        if host_ext is not None:
            if cuda:
                if not hasattr(self, "_side"):
                    self._side = torch.cuda.Stream(device=self.device)
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(self.device))
                self._side.wait_event(ev)
                with torch.cuda.stream(self._side):
                    host_ext.copy_(self.slab, non_blocking=True)
            else:
                host_ext.copy_(self.slab)
        self.merkle()
        root = self.finish()
        if host_tree is not None and self.rank == 0:
            host_tree.copy_(self.tree, non_blocking=True)
        if cuda:
            torch.cuda.current_stream(self.device).synchronize()
            if host_ext is not None:
                self._side.synchronize()
        return root


class _CudaBuffer:
    """A raw device allocation seen through __cuda_array_interface__, so that torch can view it as a tensor."""

    def __init__(self, ptr, nwords):
        self.__cuda_array_interface__ = {"shape": (nwords,), "typestr": "<i8", "data": (ptr, False), "version": 2}


def _tensor_from_ptr(ptr, nwords, device):
    return torch.as_tensor(_CudaBuffer(ptr, nwords), device=device)


def to_numpy_u64(t):
    return t.detach().cpu().numpy().view(np.uint64)
