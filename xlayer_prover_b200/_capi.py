"""ctypes binding of the C ABI declared in include/glprover.h.

`load()` opens the in-tree libglprover.so (built by `python -m xlayer_prover_b200.build` / __graft_entry__.build()).
There is no CPU implementation behind it: if the library is missing, or no CUDA device is usable,
this raises -- it never falls back to anything.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_NAME = "libglprover.so"
P = 0xFFFFFFFF00000001

GLP_OK, GLP_E_ARG, GLP_E_DOMAIN, GLP_E_CUDA, GLP_E_OOM = 0, -1, -2, -3, -4

u64, u32, i32, vp = C.c_uint64, C.c_uint32, C.c_int, C.c_void_p
u64p = C.POINTER(C.c_uint64)

# name -> (restype, argtypes); every symbol include/glprover.h declares
SIGNATURES = {
    "glp_init": (i32, [i32, C.POINTER(vp)]),
    "glp_destroy": (None, [vp]),
    "glp_last_error": (C.c_char_p, []),
    "glp_set_stream": (i32, [vp, vp]),
    "glp_synchronize": (i32, [vp]),
    "glp_launch_count": (u64, [vp]),
    "glp_leaf_redo_rows": (u64, [vp]),
    "glp_ntt_retry_count": (u64, [vp]),
    "glp_set_scratch_limit": (i32, [vp, u64]),
    "glp_release_scratch": (i32, [vp]),
    "glp_device_alloc": (i32, [vp, u64, C.POINTER(vp)]),
    "glp_device_free": (i32, [vp, vp]),
    "glp_host_alloc": (i32, [vp, u64, C.POINTER(vp)]),
    "glp_host_free": (i32, [vp, vp]),
    "glp_host_register": (i32, [vp, vp, u64]),
    "glp_host_unregister": (i32, [vp, vp]),
    "glp_memcpy": (i32, [vp, vp, vp, u64]),
    "glp_mem_info": (i32, [vp, u64p, u64p]),
    "glp_ntt": (i32, [vp, vp, vp, u64, u64, i32, i32]),
    "glp_ntt_live": (i32, [vp, vp, vp, u64, u64, i32, i32, u64]),
    "glp_extend_pol": (i32, [vp, vp, vp, u64, u64, u64]),
    "glp_extend_pol_pitched": (i32, [vp, vp, u64, vp, u64, u64, u64, u64]),
    "glp_extend_pol_scatter": (i32, [vp, C.POINTER(vp), u32, u64, u64, vp, u64, u64, u64, u64]),
    "glp_ipc_export": (i32, [vp, vp, C.c_char_p]),
    "glp_ipc_import": (i32, [vp, C.c_char_p, C.POINTER(vp)]),
    "glp_ipc_close": (i32, [vp, vp]),
    "glp_reverse_permutation": (i32, [vp, vp, vp, u64, u64, u64, u64, u64]),
    "glp_root": (u64, [u32, u64]),
    "glp_hash_full_result": (i32, [vp, vp, vp, u64]),
    "glp_linear_hash_rows": (i32, [vp, vp, vp, u64, u64, u64]),
    "glp_linear_hash_absorb": (i32, [vp, vp, vp, u64, u64, u64, i32]),
    "glp_merkletree": (i32, [vp, vp, vp, u64, u64, u64]),
    "glp_merkletree_batch": (i32, [vp, vp, vp, u64, u64, u64, u64]),
    "glp_merkle_levels": (i32, [vp, vp, u64]),
    "glp_tree_num_elements": (u64, [u64]),
    "glp_merkle_group_proofs": (i32, [vp, vp, vp, vp, u64, u64, vp, u64]),
    "glp_merkle_group_proof": (i32, [vp, vp, vp, vp, u64, u64, u64]),
    "glp_merkle_group_proofs_static": (i32, [vp, vp, vp, vp, u64, u64, vp, u64]),
    "glp_commit": (i32, [vp, vp, vp, vp, vp, u64, u64, u64]),
    "glp_quotient_commit": (i32, [vp, vp, vp, vp, vp, u64, u64, u64, u64]),
    "glp_const_tree_num_elements": (u64, [u64, u64]),
    "glp_const_tree": (i32, [vp, vp, vp, u64, u64, u64]),
    "glp_fri_fold": (i32, [vp, vp, vp, u32, u32, vp, u64]),
    "glp_fri_group_tree": (i32, [vp, vp, vp, vp, vp, u32, u32]),
    "glp_synth_fill": (i32, [vp, vp, u64, u64, u64, u64, u64, u64]),
    "glp_fib_fill": (i32, [vp, vp, u64, u64]),
    "glp_checksum": (i32, [vp, vp, u64, u64p]),
    "glp_field_selftest": (i32, [vp, vp, vp, u64, vp]),
    "glp_set_profiling": (i32, [vp, i32]),
    "glp_get_profile": (i32, [vp, C.POINTER(C.c_float), C.POINTER(u32)]),
    "glp_measure_int_peak": (i32, [vp, C.POINTER(C.c_double)]),
    "glp_measure_pipe_peaks": (i32, [vp, C.POINTER(C.c_double)]),
    "glp_place_subtree": (i32, [vp, vp, vp, u64, u32, u32]),
    "glp_multi_init": (i32, [i32, C.POINTER(i32), C.POINTER(vp)]),
    "glp_multi_destroy": (None, [vp]),
    "glp_multi_device_count": (i32, [vp]),
    "glp_multi_ctx": (vp, [vp, i32]),
    "glp_multi_extend_pol": (i32, [vp, vp, vp, u64, u64, u64]),
    "glp_multi_merkletree": (i32, [vp, vp, vp, u64, u64, u64]),
    "glp_multi_commit": (i32, [vp, vp, vp, vp, vp, u64, u64, u64]),
}


class GlpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("glprover error %d: %s" % (code, msg))
        self.code = code


class GlpLib:
    """The shared library with typed entry points."""

    def __init__(self, path):
        if not os.path.exists(path):
            raise FileNotFoundError(
                "%s not found: build it with `python -m xlayer_prover_b200.build` "
                "(nvcc, sm_100a).  There is no CPU fallback." % path)
        self.path = path
        self.cdll = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(self.cdll, name)      # AttributeError if the ABI and the library disagree
            f.restype, f.argtypes = res, args
            setattr(self, name, f)

    def last_error(self):
        return (self.glp_last_error() or b"").decode()

    def check(self, rc):
        if rc != GLP_OK:
            raise GlpError(rc, self.last_error())


_lib = None


def load():
    global _lib
    if _lib is None:
        _lib = GlpLib(os.path.join(HERE, LIB_NAME))
    return _lib


def ptr_of(a):
    """Address of a numpy array, a torch tensor (host or CUDA), a raw int address, or None."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        if a.dtype != np.uint64 or not a.flags["C_CONTIGUOUS"]:
            raise TypeError("expected a C-contiguous uint64 array")
        return a.ctypes.data
    if hasattr(a, "data_ptr"):            # torch.Tensor (int64 / uint64 storage, same bytes)
        if a.element_size() != 8 or not a.is_contiguous():
            raise TypeError("expected a contiguous 8-byte-element tensor")
        return a.data_ptr()
    raise TypeError("unsupported buffer type %r" % type(a))


class Context:
    """One glp_ctx: one device, one stream (include/glprover.h)."""

    def __init__(self, device=0, lib=None):
        self.lib = lib or load()
        h = vp()
        self.lib.check(self.lib.glp_init(device, C.byref(h)))
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.glp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- plumbing
    def set_stream(self, raw_stream): self.lib.check(self.lib.glp_set_stream(self.h, raw_stream))
    def synchronize(self): self.lib.check(self.lib.glp_synchronize(self.h))
    def launch_count(self): return self.lib.glp_launch_count(self.h)
    def leaf_redo_rows(self): return int(self.lib.glp_leaf_redo_rows(self.h))
    def ntt_retry_count(self): return int(self.lib.glp_ntt_retry_count(self.h))
    def set_scratch_limit(self, nbytes): self.lib.check(self.lib.glp_set_scratch_limit(self.h, nbytes))

    def release_scratch(self): self.lib.check(self.lib.glp_release_scratch(self.h))

    def device_alloc(self, nbytes):
        p = vp()
        self.lib.check(self.lib.glp_device_alloc(self.h, nbytes, C.byref(p)))
        return p.value

    def device_free(self, p): self.lib.check(self.lib.glp_device_free(self.h, p))

    def host_alloc(self, nbytes):
        p = vp()
        self.lib.check(self.lib.glp_host_alloc(self.h, nbytes, C.byref(p)))
        return p.value

    def host_free(self, p): self.lib.check(self.lib.glp_host_free(self.h, p))
    def host_register(self, a):
        """pin an existing numpy array (glp_host_register); pair with host_unregister"""
        self.lib.check(self.lib.glp_host_register(self.h, ptr_of(a), a.nbytes))

    def host_unregister(self, a): self.lib.check(self.lib.glp_host_unregister(self.h, ptr_of(a)))

    def memcpy(self, dst, src, nbytes): self.lib.check(self.lib.glp_memcpy(self.h, ptr_of(dst), ptr_of(src), nbytes))

    def mem_info(self):
        f, t = u64(), u64()
        self.lib.check(self.lib.glp_mem_info(self.h, C.byref(f), C.byref(t)))
        return f.value, t.value

    def to_device(self, a):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        p = self.device_alloc(max(a.nbytes, 8))
        self.memcpy(p, a, a.nbytes)
        return p

    def to_host(self, p, shape):
        out = np.empty(shape, dtype=np.uint64)
        self.memcpy(out, p, out.nbytes)
        return out

    # ---- the path
    def ntt(self, dst, src, size, ncols=1, inverse=False, extend=False, live_rows=None):
        """live_rows: rows >= live_rows of src are read as zero (an NTT_Goldilocks object built with extension > 1)"""
        if live_rows is None:
            self.lib.check(self.lib.glp_ntt(self.h, ptr_of(dst), ptr_of(src), size, ncols, int(inverse), int(extend)))
        else:
            self.lib.check(self.lib.glp_ntt_live(self.h, ptr_of(dst), ptr_of(src), size, ncols, int(inverse), int(extend), live_rows))

    def fib_fill(self, dst, nrows, ncols):
        """the reference's Fibonacci test matrix (bench.cpp:410-427), generated on the device"""
        self.lib.check(self.lib.glp_fib_fill(self.h, ptr_of(dst), nrows, ncols))

    def extend_pol(self, out, inp, n_ext, n, ncols):
        self.lib.check(self.lib.glp_extend_pol(self.h, ptr_of(out), ptr_of(inp), n_ext, n, ncols))

    def extend_pol_pitched(self, out, out_pitch, inp, in_pitch, n_ext, n, ncols):
        self.lib.check(self.lib.glp_extend_pol_pitched(self.h, ptr_of(out), out_pitch, ptr_of(inp), in_pitch, n_ext, n, ncols))

    def extend_pol_scatter(self, parts, part_pitch, part_col0, inp, in_pitch, n_ext, n, ncols):
        """parts: list of slab base addresses (ints or tensors), one per rank."""
        arr = (vp * len(parts))(*[ptr_of(x) for x in parts])
        self.lib.check(self.lib.glp_extend_pol_scatter(self.h, arr, len(parts), part_pitch, part_col0, ptr_of(inp), in_pitch, n_ext, n, ncols))

    def ipc_export(self, devptr):
        buf = C.create_string_buffer(64)
        self.lib.check(self.lib.glp_ipc_export(self.h, ptr_of(devptr), buf))
        return buf.raw

    def ipc_import(self, handle):
        p = vp()
        self.lib.check(self.lib.glp_ipc_import(self.h, handle, C.byref(p)))
        return p.value

    def ipc_close(self, devptr): self.lib.check(self.lib.glp_ipc_close(self.h, devptr))

    def reverse_permutation(self, dst, src, size, offset_cols, ncols, ncols_all, extension=1):
        self.lib.check(self.lib.glp_reverse_permutation(self.h, ptr_of(dst), ptr_of(src), size, offset_cols, ncols, ncols_all, extension))

    def hash_full_result(self, out, inp, n=1):
        self.lib.check(self.lib.glp_hash_full_result(self.h, ptr_of(out), ptr_of(inp), n))

    def linear_hash_rows(self, out4, inp, ncols, pitch, nrows):
        self.lib.check(self.lib.glp_linear_hash_rows(self.h, ptr_of(out4), ptr_of(inp), ncols, pitch, nrows))

    def linear_hash_absorb(self, state4, inp, ncols, pitch, nrows, first):
        self.lib.check(self.lib.glp_linear_hash_absorb(self.h, ptr_of(state4), ptr_of(inp), ncols, pitch, nrows, int(first)))

    def merkletree(self, tree, inp, ncols, nrows, dim=1):
        self.lib.check(self.lib.glp_merkletree(self.h, ptr_of(tree), ptr_of(inp), ncols, nrows, dim))

    def merkletree_batch(self, tree, inp, ncols, nrows, batch_size, dim=1):
        self.lib.check(self.lib.glp_merkletree_batch(self.h, ptr_of(tree), ptr_of(inp), ncols, nrows, batch_size, dim))

    def merkle_levels(self, level, n):
        self.lib.check(self.lib.glp_merkle_levels(self.h, ptr_of(level), n))

    def merkle_group_proof(self, proof, source, nodes, width, height, idx):
        self.lib.check(self.lib.glp_merkle_group_proof(self.h, ptr_of(proof), ptr_of(source), ptr_of(nodes), width, height, idx))

    def place_subtree(self, tree, subtree, sub_rows, nparts, part):
        """queue the copies of this rank's subtree levels into their slices of the full tree (stream ordered, not synchronised)"""
        self.lib.check(self.lib.glp_place_subtree(self.h, ptr_of(tree), ptr_of(subtree), sub_rows, nparts, part))

    def merkle_group_proofs(self, proofs, source, nodes, width, height, idx):
        """getGroupProof for every index of `idx` (a host uint64 array) in one launch / one download"""
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        self.lib.check(self.lib.glp_merkle_group_proofs(self.h, ptr_of(proofs), ptr_of(source), ptr_of(nodes), width, height, ptr_of(idx), idx.size))

    def merkle_group_proofs_static(self, proofs, source, nodes, width, height, idx):
        """merkle_group_proofs with the index staging in a fixed device array: no allocation; device buffers, at most
        GLP_GROUP_PROOFS_STATIC_MAX (4096) indices"""
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        self.lib.check(self.lib.glp_merkle_group_proofs_static(self.h, ptr_of(proofs), ptr_of(source), ptr_of(nodes), width, height, ptr_of(idx),
                                                               idx.size))

    def commit(self, out_ext, tree, inp, n, n_ext, ncols):
        root = np.zeros(4, dtype=np.uint64)
        self.lib.check(self.lib.glp_commit(self.h, ptr_of(out_ext), ptr_of(tree), ptr_of(root), ptr_of(inp), n, n_ext, ncols))
        return root

    def quotient_commit(self, cm4, tree, q, n, n_ext, qdim, qdeg):
        """Starks::genProof step 4 (starks.cpp:261-291): cm4 = NTT(split(INTT(q))), tree over cm4; returns the root."""
        root = np.zeros(4, dtype=np.uint64)
        self.lib.check(self.lib.glp_quotient_commit(self.h, ptr_of(cm4), ptr_of(tree), ptr_of(root), ptr_of(q), n, n_ext, qdim, qdeg))
        return root

    def const_tree(self, const_tree, const_pols, n, n_ext, npols):
        """bctree's *.consttree layout [npols, n_ext, extended pols, nodes] (build_const_tree.cpp:555-575)."""
        self.lib.check(self.lib.glp_const_tree(self.h, ptr_of(const_tree), ptr_of(const_pols), n, n_ext, npols))

    def fri_fold(self, out, pol, pol_bits, step_bits, special_x, shift_inv):
        """one folding step of FRIProve::prove (friProve.cpp:22-108): 2^pol_bits -> 2^step_bits extension elements"""
        x = (u64 * 3)(*[int(v) for v in special_x])
        self.lib.check(self.lib.glp_fri_fold(self.h, ptr_of(out), ptr_of(pol), pol_bits, step_bits, x, int(shift_inv)))

    def fri_group_tree(self, aux, tree, pol, pol_bits, next_bits):
        """getTransposed + MerkleTreeGL of a FRI layer (friProve.cpp:110-123); returns the root"""
        root = np.zeros(4, dtype=np.uint64)
        self.lib.check(self.lib.glp_fri_group_tree(self.h, ptr_of(aux), ptr_of(tree), ptr_of(root), ptr_of(pol), pol_bits, next_bits))
        return root

    def synth_fill(self, dst, nrows, ncols, seed, row0=0, col0=0, ncols_total=None):
        self.lib.check(self.lib.glp_synth_fill(self.h, ptr_of(dst), nrows, ncols, row0, col0,
                                                ncols if ncols_total is None else ncols_total, seed))

    def checksum(self, src, nwords):
        out = (u64 * 2)()
        self.lib.check(self.lib.glp_checksum(self.h, ptr_of(src), nwords, out))
        return int(out[0]), int(out[1])

    def field_selftest(self, a, b):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        b = np.ascontiguousarray(b, dtype=np.uint64)
        out = np.zeros((a.size, 8), dtype=np.uint64)
        self.lib.check(self.lib.glp_field_selftest(self.h, ptr_of(a), ptr_of(b), a.size, ptr_of(out)))
        return out


    def set_profiling(self, on): self.lib.check(self.lib.glp_set_profiling(self.h, int(on)))

    def get_profile(self):
        """{stage: (ms, intervals)} accumulated since the last call; synchronises the stream."""
        ms, cnt = (C.c_float * 4)(), (u32 * 4)()
        self.lib.check(self.lib.glp_get_profile(self.h, ms, cnt))
        names = ("intt", "ntt_ext", "leaf_hash", "tree_levels")
        return {n: (float(ms[i]), int(cnt[i])) for i, n in enumerate(names)}

    def measure_int_peak(self):
        out = (C.c_double * 3)()
        self.lib.check(self.lib.glp_measure_int_peak(self.h, out))
        return {"imad_lane_ops_per_s": out[0], "iadd3_lane_ops_per_s": out[1], "interleaved_lane_ops_per_s": out[2]}

    def measure_pipe_peaks(self):
        """lane-ops/s of IMAD, IADD3, both interleaved, IMAD.WIDE (64-bit accumulate) and DFMA on this device"""
        out = (C.c_double * 5)()
        self.lib.check(self.lib.glp_measure_pipe_peaks(self.h, out))
        return {"imad_lane_ops_per_s": out[0], "iadd3_lane_ops_per_s": out[1], "interleaved_lane_ops_per_s": out[2],
                "imad_wide_lane_ops_per_s": out[3], "dfma_lane_ops_per_s": out[4]}


class MultiContext:
    """One glp_multi: G devices driven from this process (include/glprover.h, csrc/multi.cuh).  Host buffers only."""

    def __init__(self, devices, lib=None):
        self.lib = lib or load()
        devs = (i32 * len(devices))(*devices)
        h = vp()
        self.lib.check(self.lib.glp_multi_init(len(devices), devs, C.byref(h)))
        self.h = h
        self.devices = list(devices)

    def close(self):
        if getattr(self, "h", None):
            self.lib.glp_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def device_count(self): return int(self.lib.glp_multi_device_count(self.h))

    def extend_pol(self, out, inp, n_ext, n, ncols):
        self.lib.check(self.lib.glp_multi_extend_pol(self.h, ptr_of(out), ptr_of(inp), n_ext, n, ncols))

    def merkletree(self, tree, inp, ncols, nrows, dim=1):
        self.lib.check(self.lib.glp_multi_merkletree(self.h, ptr_of(tree), ptr_of(inp), ncols, nrows, dim))

    def commit(self, out_ext, tree, inp, n, n_ext, ncols):
        root = np.zeros(4, dtype=np.uint64)
        self.lib.check(self.lib.glp_multi_commit(self.h, ptr_of(out_ext), ptr_of(tree), ptr_of(root), ptr_of(inp), n, n_ext, ncols))
        return root


def tree_num_elements(nrows):
    """MerklehashGoldilocks::getTreeNumElements (merklehash_goldilocks.hpp:24-27)."""
    return (2 * nrows - 1) * 4 if nrows else 0


def splitmix(seed, idx):
    """Host twin of the device generator in csrc/aux_kernels.cuh (numpy, vectorised)."""
    idx = np.asarray(idx, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = (np.uint64(seed) ^ idx) + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
        return np.where(z >= np.uint64(P), z - np.uint64(P), z)


def synth_matrix(nrows, ncols, seed, row0=0, col0=0, ncols_total=None):
    nt = ncols if ncols_total is None else ncols_total
    r = np.arange(row0, row0 + nrows, dtype=np.uint64)[:, None]
    c = np.arange(col0, col0 + ncols, dtype=np.uint64)[None, :]
    with np.errstate(over="ignore"):
        return splitmix(seed, r * np.uint64(nt) + c)
