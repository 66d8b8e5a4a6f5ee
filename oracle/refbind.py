"""TEST INFRASTRUCTURE ONLY -- ctypes binding over oracle/_ref/libglref*.so (the unmodified
reference compiled by oracle/Makefile) and oracle/libgloracle.so (the plain-C restatement).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package (xlayer_prover_b200) never does.
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
P = 0xFFFFFFFF00000001
u64p = C.POINTER(C.c_uint64)


def _ptr(a):
    if a is None:
        return None
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(u64p)


class _Lib:
    """Common face of the reference build (prefix ref_) and the C port (prefix orc_)."""

    def __init__(self, path, prefix):
        self.path = path
        self.lib = C.CDLL(path)
        self.prefix = prefix
        L = self.lib
        g = lambda n: getattr(L, prefix + n)
        u64, i32, u32 = C.c_uint64, C.c_int, C.c_uint32
        sig = {
            "add": (u64, [u64, u64]), "sub": (u64, [u64, u64]), "mul": (u64, [u64, u64]),
            "to_u64": (u64, [u64]), "w": (u64, [u32]), "shift": (u64, []),
            "canonicalize": (None, [u64p, u64]),
            "ntt": (i32, [u64p, u64p, u64, u64, u64, u64, i32, i32]),
            "extend_pol": (i32, [u64p, u64p, u64, u64, u64, u64p, u64, u64]),
            "hash_full_result": (None, [u64p, u64p]),
            "hash_full_result_seq": (None, [u64p, u64p]),
            "linear_hash": (None, [u64p, u64p, u64]),
            "linear_hash_rows": (None, [u64p, u64p, u64, u64]),
            "merkletree": (None, [u64p, u64p, u64, u64, i32, u64]),
            "merkletree_batch": (None, [u64p, u64p, u64, u64, u64, i32, u64]),
            "tree_num_elements": (u64, [u64]),
            "max_threads": (i32, []), "set_threads": (None, [i32]),
            "commit": (i32, [u64p, u64p, u64p, u64, u64, u64, u64p, C.POINTER(C.c_double)]),
        }
        for name, (res, args) in sig.items():
            f = g(name)
            f.restype, f.argtypes = res, args
            setattr(self, "_" + name, f)

    # ---- scalars
    def add(self, a, b): return self._add(a, b)
    def sub(self, a, b): return self._sub(a, b)
    def mul(self, a, b): return self._mul(a, b)
    def to_u64(self, a): return self._to_u64(a)
    def w(self, k): return self._w(k)
    def shift(self): return self._shift()
    def max_threads(self): return self._max_threads()
    def set_threads(self, n): self._set_threads(n)

    def canonical(self, a):
        out = np.ascontiguousarray(a, dtype=np.uint64).copy()
        self._canonicalize(_ptr(out.reshape(-1)), out.size)
        return out

    # ---- NTT
    def ntt(self, src, inverse=False, extend=False, nphase=3, nblock=1):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        if src.size == 0:
            return src.copy()
        a = src.reshape(src.shape[0], -1) if src.ndim > 1 else src.reshape(-1, 1)
        size, ncols = a.shape
        dst = np.zeros_like(a)
        if size and ncols:
            rc = self._ntt(_ptr(dst), _ptr(a.copy()), size, ncols, nphase, nblock, int(inverse), int(extend))
            if rc:
                raise ValueError("reference ntt rc=%d" % rc)
        return self.canonical(dst).reshape(src.shape)

    def extend_pol(self, inp, n_ext, nphase=3, nblock=1):
        inp = np.ascontiguousarray(inp, dtype=np.uint64)
        n, ncols = inp.shape
        out = np.zeros((n_ext, ncols), dtype=np.uint64)
        buf = np.zeros((n_ext, ncols), dtype=np.uint64)
        self._extend_pol(_ptr(out), _ptr(inp.copy()), n_ext, n, ncols, _ptr(buf), nphase, nblock)
        return self.canonical(out)

    # ---- Poseidon
    def hash_full_result(self, x, seq=False):
        x = np.ascontiguousarray(x, dtype=np.uint64)
        out = np.zeros(12, dtype=np.uint64)
        (self._hash_full_result_seq if seq else self._hash_full_result)(_ptr(out), _ptr(x))
        return self.canonical(out)

    def linear_hash(self, x):
        x = np.ascontiguousarray(x, dtype=np.uint64).reshape(-1)
        out = np.zeros(4, dtype=np.uint64)
        self._linear_hash(_ptr(out), _ptr(x) if x.size else None, x.size)
        return self.canonical(out)

    def linear_hash_rows(self, x):
        x = np.ascontiguousarray(x, dtype=np.uint64)
        nrows, ncols = x.shape
        out = np.zeros((nrows, 4), dtype=np.uint64)
        self._linear_hash_rows(_ptr(out), _ptr(x) if x.size else None, ncols, nrows)
        return self.canonical(out)

    def merkletree(self, x, nrows=None, ncols=None, nthreads=0):
        if x is not None:
            x = np.ascontiguousarray(x, dtype=np.uint64)
            nrows, ncols = x.shape
        tree = np.zeros(self._tree_num_elements(nrows) if nrows else 0, dtype=np.uint64)
        if nrows:
            self._merkletree(_ptr(tree), _ptr(x) if (x is not None and x.size) else None, ncols, nrows, nthreads, 1)
        return self.canonical(tree)

    def merkletree_batch(self, x, batch, nthreads=0):
        x = np.ascontiguousarray(x, dtype=np.uint64)
        nrows, ncols = x.shape
        tree = np.zeros(self._tree_num_elements(nrows), dtype=np.uint64)
        self._merkletree_batch(_ptr(tree), _ptr(x), ncols, nrows, batch, nthreads, 1)
        return self.canonical(tree)

    def commit(self, inp, n_ext):
        """extendPol + merkletree as Starks::genProof step 1 runs them; returns (ext, tree, (t_lde, t_merkle))."""
        inp = np.ascontiguousarray(inp, dtype=np.uint64)
        n, ncols = inp.shape
        out = np.zeros((n_ext, ncols), dtype=np.uint64)
        buf = np.zeros((n_ext, ncols), dtype=np.uint64)
        tree = np.zeros(self._tree_num_elements(n_ext), dtype=np.uint64)
        secs = (C.c_double * 2)()
        self._commit(_ptr(out), _ptr(tree), _ptr(inp), n, n_ext, ncols, _ptr(buf), secs)
        return out, tree, (secs[0], secs[1])


    def timed_parts(self, inp, n_ext, leaf_rows):
        """The three pieces bench.py's CPU arm times, each one call of the library's own entry point on preallocated
        buffers (perf_counter around the foreign call): extendPol of `inp` to n_ext rows; merkletree of the first
        `leaf_rows` rows of the extension (leaf hashes + the levels above them); merkletree of a 4-column matrix of
        `leaf_rows` rows (<= 4 columns are copied into the leaves unhashed, so this is the level loop alone).
        Returns (t_extend_pol, t_tree_sample, t_levels_sample)."""
        import time
        inp = np.ascontiguousarray(inp, dtype=np.uint64)
        n, ncols = inp.shape
        out = np.zeros((n_ext, ncols), dtype=np.uint64)
        buf = np.zeros((n_ext, ncols), dtype=np.uint64)
        src = inp.copy()                                   # the reference's INTT works in place on its input when it can
        t0 = time.perf_counter()
        self._extend_pol(_ptr(out), _ptr(src), n_ext, n, ncols, _ptr(buf), 3, 1)
        t_lde = time.perf_counter() - t0
        del buf, src
        tree = np.zeros(self._tree_num_elements(leaf_rows), dtype=np.uint64)
        t0 = time.perf_counter()
        self._merkletree(_ptr(tree), _ptr(out), ncols, leaf_rows, 0, 1)
        t_tree = time.perf_counter() - t0
        four = np.ascontiguousarray(out[:leaf_rows, :4]) if ncols >= 4 else np.zeros((leaf_rows, 4), dtype=np.uint64)
        t0 = time.perf_counter()
        self._merkletree(_ptr(tree), _ptr(four), 4, leaf_rows, 0, 1)
        t_levels = time.perf_counter() - t0
        return t_lde, t_tree, t_levels


class RefLib(_Lib):
    """The compiled reference, plus the entry points only it has (MerkleTreeGL, root, ...)."""

    def __init__(self, path):
        super().__init__(path, "ref_")
        L = self.lib
        u64, vp = C.c_uint64, C.c_void_p
        L.ref_is_avx512.restype = C.c_int
        L.ref_root.restype, L.ref_root.argtypes = u64, [u64, C.c_uint32, u64]
        L.ref_reverse_permutation.restype, L.ref_reverse_permutation.argtypes = None, [u64p, u64p, u64, u64, u64, u64]
        L.ref_ntt_maxsize.restype, L.ref_ntt_maxsize.argtypes = C.c_int, [u64p, u64p, u64, u64, u64, C.c_int]
        L.ref_merkletree_seq.restype, L.ref_merkletree_seq.argtypes = None, [u64p, u64p, u64, u64, C.c_int, u64]
        L.ref_mt_new.restype, L.ref_mt_new.argtypes = vp, [u64, u64, u64p]
        L.ref_mt_map.restype, L.ref_mt_map.argtypes = vp, [u64p]
        L.ref_mt_free.restype, L.ref_mt_free.argtypes = None, [vp]
        L.ref_mt_merkelize.restype, L.ref_mt_merkelize.argtypes = None, [vp]
        L.ref_mt_nodes.restype, L.ref_mt_nodes.argtypes = u64p, [vp]
        L.ref_mt_num_elements.restype, L.ref_mt_num_elements.argtypes = u64, [vp]
        L.ref_mt_proof_size.restype, L.ref_mt_proof_size.argtypes = u64, [vp]
        L.ref_mt_get_root.restype, L.ref_mt_get_root.argtypes = None, [vp, u64p]
        L.ref_mt_get_group_proof.restype, L.ref_mt_get_group_proof.argtypes = None, [vp, u64p, u64]

        L.ref_quotient_commit.restype, L.ref_quotient_commit.argtypes = C.c_int, [u64p, u64p, u64p, u64, u64, u64, u64]
        L.ref_fri_group_tree.restype, L.ref_fri_group_tree.argtypes = C.c_int, [u64p, u64p, u64p, u64, u64]
        L.ref_tr_new.restype, L.ref_tr_new.argtypes = vp, []
        L.ref_tr_free.restype, L.ref_tr_free.argtypes = None, [vp]
        L.ref_tr_put.restype, L.ref_tr_put.argtypes = None, [vp, u64p, u64]
        L.ref_tr_get_field.restype, L.ref_tr_get_field.argtypes = None, [vp, u64p]
        L.ref_tr_get_fields1.restype, L.ref_tr_get_fields1.argtypes = u64, [vp]
        L.ref_tr_get_permutations.restype, L.ref_tr_get_permutations.argtypes = None, [vp, u64p, u64, u64]

    def quotient_commit(self, q, n, qdeg):
        """Starks::genProof step 4 (starks.cpp:261-291) on q [n_ext x qdim]; returns (cm4, tree), canonical."""
        q = np.ascontiguousarray(q, dtype=np.uint64)
        n_ext, qdim = q.shape
        cm4 = np.zeros((n_ext, qdim * qdeg), dtype=np.uint64)
        tree = np.zeros(self._tree_num_elements(n_ext), dtype=np.uint64)
        self.lib.ref_quotient_commit(_ptr(cm4), _ptr(tree), _ptr(q.copy()), n, n_ext, qdim, qdeg)
        return self.canonical(cm4), self.canonical(tree)

    # ---- the reference's constant-tree builder and FRI prover, compiled from their own sources (oracle/_ref/libbctree.so,
    # libfriref.so; oracle/Makefile): no arithmetic of theirs is restated on this side
    def _bct(self):
        if getattr(self, "_bctlib", None) is None:
            L = C.CDLL(os.path.join(os.path.dirname(self.path), "libbctree.so"))
            L.ref_bctree_interpolate.restype, L.ref_bctree_interpolate.argtypes = C.c_int, [u64p, u64p, C.c_uint64, C.c_uint64, C.c_uint64]
            L.ref_bctree_build.restype, L.ref_bctree_build.argtypes = C.c_int, [C.c_char_p] * 4
            self._bctlib = L
        return self._bctlib

    def _fri(self):
        if getattr(self, "_frilib", None) is None:
            L = C.CDLL(os.path.join(os.path.dirname(self.path), "libfriref.so"))
            u64 = C.c_uint64
            L.ref_fri_fold_step.restype, L.ref_fri_fold_step.argtypes = C.c_int, [u64p, u64p, u64, u64, u64p, u64]
            L.ref_fri_prove.restype = C.c_int
            L.ref_fri_prove.argtypes = [u64p, u64, u64p, u64, u64, C.POINTER(u64p), u64p, u64p, u64, u64p, u64p, u64p, u64p, u64, u64p]
            self._frilib = L
        return self._frilib

    def bctree_interpolate(self, const_pols, n_ext):
        """bctree's own `interpolate` (tools/starkpil/bctree/build_const_tree.cpp:398-531): [n x npols] -> [n_ext x npols], canonical"""
        a = np.ascontiguousarray(const_pols, dtype=np.uint64)
        n, npols = a.shape
        out = np.zeros((n_ext, npols), dtype=np.uint64)
        with self._at_least_two_threads():
            self._bct().ref_bctree_interpolate(_ptr(out), _ptr(a.copy()), npols, n.bit_length() - 1, n_ext.bit_length() - 1)
        return self.canonical(out)

    def _at_least_two_threads(self):
        """bctree divides by omp_get_max_threads() / 2 (build_const_tree.cpp:408,370): with one OpenMP thread that is a division by zero"""
        import contextlib

        @contextlib.contextmanager
        def cm():
            t = self.max_threads()
            if t < 2:
                self.set_threads(2)
            try:
                yield
            finally:
                if t < 2:
                    self.set_threads(t)
        return cm()

    def bctree_build(self, const_file, stark_struct_file, const_tree_file, ver_key_file):
        """buildConstTree (build_const_tree.cpp:533-600): the files bctree reads and writes"""
        with self._at_least_two_threads():
            self._bct().ref_bctree_build(*(os.fsencode(f) for f in (const_file, stark_struct_file, const_tree_file, ver_key_file)))

    def const_tree(self, const_pols, n_ext):
        """bctree's *.consttree words [nPols, nExt, extended pols, nodes] (build_const_tree.cpp:555-575), canonical: bctree's
        interpolate, then the reference's merkletree, as buildConstTree strings them together (:545-575)."""
        a = np.ascontiguousarray(const_pols, dtype=np.uint64)
        n, npols = a.shape
        ext = self.bctree_interpolate(a, n_ext)
        return np.concatenate([np.array([npols, n_ext], dtype=np.uint64), ext.reshape(-1), self.merkletree(ext)])

    def fri_fold(self, pol, pol_bits, step_bits, special_x, shift_inv):
        """one folding step with the reference's own INTT, FRIProve::polMulAxi and FRIProve::evalPol (friProve.cpp:93-105,192-217)"""
        pol = np.ascontiguousarray(pol, dtype=np.uint64)
        out = np.zeros((1 << step_bits, 3), dtype=np.uint64)
        x = np.ascontiguousarray(special_x, dtype=np.uint64)
        self._fri().ref_fri_fold_step(_ptr(out), _ptr(pol.copy()), pol_bits, step_bits, _ptr(x), int(shift_inv))
        return self.canonical(out)

    def fri_prove(self, fri_pol, pol_bits, step_bits, nqueries, trees_src, seed):
        """FRIProve::prove itself (friProve.cpp:5-190) on friPol [2^pol_bits x 3] and the five stage trees whose sources are
        trees_src[i] [2^pol_bits x w_i]; the transcript holds `seed` when it starts.  Returns what it left in the FRIProof:
        roots of the layer trees, the last polynomial, and the query proofs flattened in the order queryPol pushed them."""
        pol = np.ascontiguousarray(fri_pol, dtype=np.uint64).copy()
        srcs = [np.ascontiguousarray(t, dtype=np.uint64) for t in trees_src]
        assert len(srcs) == 5 and all(t.shape[0] == 1 << pol_bits for t in srcs)
        widths = np.array([t.shape[1] for t in srcs], dtype=np.uint64)
        ptrs = (u64p * 5)(*[_ptr(t) for t in srcs])
        steps = np.array(step_bits, dtype=np.uint64)
        seed = np.ascontiguousarray(seed, dtype=np.uint64)
        roots = np.zeros((max(0, len(step_bits) - 1), 4), dtype=np.uint64)
        final = np.zeros((1 << step_bits[-1], 3), dtype=np.uint64)
        ys = np.zeros(nqueries, dtype=np.uint64)
        cap = 0
        for si, nb in enumerate(step_bits):
            if si == 0:
                cap += nqueries * sum(int(w) + 4 * pol_bits for w in widths)
            else:
                cap += nqueries * (3 * (1 << (step_bits[si - 1] - nb)) + 4 * nb)
        q = np.zeros(cap, dtype=np.uint64)
        used = C.c_uint64(0)
        rc = self._fri().ref_fri_prove(_ptr(pol), pol_bits, _ptr(steps), len(step_bits), nqueries, ptrs, _ptr(widths), _ptr(seed), seed.size,
                                       _ptr(roots.reshape(-1)) if roots.size else None, _ptr(final.reshape(-1)), _ptr(ys), _ptr(q), cap, C.byref(used))
        assert rc == 0 and used.value == cap, (rc, used.value, cap)
        return {"roots": roots, "final": final, "queries": q}

    def fri_group_tree(self, pol, next_bits):
        """getTransposed + MerkleTreeGL of a FRI layer (friProve.cpp:110-123,252-270); returns (aux, tree), canonical"""
        pol = np.ascontiguousarray(pol, dtype=np.uint64)
        n = pol.shape[0]
        aux = np.zeros((1 << next_bits, 3 * (n >> next_bits)), dtype=np.uint64)
        tree = np.zeros(self._tree_num_elements(1 << next_bits), dtype=np.uint64)
        self.lib.ref_fri_group_tree(_ptr(aux), _ptr(tree), _ptr(pol.copy()), n, next_bits)
        return self.canonical(aux), self.canonical(tree)

    def transcript(self):
        """The reference's Transcript object behind a small Python face (put / get_field / get_fields1 / get_permutations)."""
        L = self.lib

        class _T:
            def __init__(s): s.h = L.ref_tr_new()
            def __del__(s): L.ref_tr_free(s.h)
            def put(s, a):
                a = np.ascontiguousarray(a, dtype=np.uint64).reshape(-1)
                L.ref_tr_put(s.h, _ptr(a), a.size)
            def get_fields1(s): return int(L.ref_tr_get_fields1(s.h))
            def get_field(s):
                o = np.zeros(3, dtype=np.uint64)
                L.ref_tr_get_field(s.h, _ptr(o))
                return [int(x) % P for x in o]
            def get_permutations(s, n, nbits):
                o = np.zeros(n, dtype=np.uint64)
                L.ref_tr_get_permutations(s.h, _ptr(o), n, nbits)
                return [int(x) for x in o]
        return _T()

    def is_avx512(self): return bool(self.lib.ref_is_avx512())
    def root(self, max_size, domain_pow, idx): return self.lib.ref_root(max_size, domain_pow, idx)

    def merkle_tree_gl(self, source):
        """MerkleTreeGL(height,width,source).merkelize(); returns (nodes, root, proof_size, get_group_proof)."""
        source = np.ascontiguousarray(source, dtype=np.uint64)
        h, w = source.shape
        L = self.lib
        mt = L.ref_mt_new(h, w, _ptr(source))
        L.ref_mt_merkelize(mt)
        n = L.ref_mt_num_elements(mt)
        nodes = np.ctypeslib.as_array(L.ref_mt_nodes(mt), shape=(n,)).copy()
        root = np.zeros(4, dtype=np.uint64)
        L.ref_mt_get_root(mt, _ptr(root))
        psize = L.ref_mt_proof_size(mt)
        proofs = {}

        def group_proof(idx):
            pr = np.zeros(w + 4 * psize, dtype=np.uint64)
            L.ref_mt_get_group_proof(mt, _ptr(pr), idx)
            return self.canonical(pr)
        return self.canonical(nodes), self.canonical(root), psize, group_proof


_cache = {}


def reference(avx512=False):
    """The compiled reference (oracle/_ref); None when it has not been built."""
    name = "libglref512.so" if avx512 else "libglref.so"
    path = os.path.join(HERE, "_ref", name)
    if path not in _cache:
        _cache[path] = RefLib(path) if os.path.exists(path) else None
    return _cache[path]


def port():
    """The plain-C restatement (oracle/gl_oracle.c); None when not built."""
    path = os.path.join(HERE, "libgloracle.so")
    if path not in _cache:
        _cache[path] = _Lib(path, "orc_") if os.path.exists(path) else None
    return _cache[path]


def best():
    """Reference when built, else the port (header of DESIGN.md: parity is pinned on the reference)."""
    return reference() or port()
