"""xlayer_prover_b200 -- B200-native STARK commitment hot path (Goldilocks LDE + Poseidon Merkle).

Host-side mirror of the two reference classes the hot path sits behind, over the C ABI of
libglprover.so (include/glprover.h):

    NTT_Goldilocks   src/goldilocks/src/ntt_goldilocks.hpp:12-174
    MerkleTreeGL     src/starkpil/merkleTree/merkleTreeGL.hpp:9-79
    PoseidonGoldilocks (static functions)  src/goldilocks/src/poseidon_goldilocks.hpp:17-67

Same names, argument order and error behaviour as the reference, so the parity tests read like the
reference's own tests (src/goldilocks/tests/tests.cpp).  Buffers are numpy uint64 arrays (host) or
torch CUDA tensors / raw device addresses (device); results are written into the caller's buffer.
The C++ drop-in headers for starks.hpp are include/ntt_goldilocks.hpp and include/merkleTreeGL.hpp.

All compute runs in the CUDA library; if it is missing or there is no GPU this package raises.
"""
import math

import numpy as np

from ._capi import (Context, GlpError, GlpLib, MultiContext, P, load, ptr_of, splitmix, synth_matrix, tree_num_elements,
                    GLP_E_ARG, GLP_E_DOMAIN, GLP_E_CUDA, GLP_E_OOM)

RATE, CAPACITY, HASH_SIZE, SPONGE_WIDTH = 8, 4, 4, 12
NUM_PHASES, NUM_BLOCKS = 3, 1

_default_ctx = {}


def default_context(device=0):
    """Lazily created per-device context (construction of the wrapper classes does no GPU work)."""
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]


class NTT_Goldilocks:
    """NTT_Goldilocks(maxDomainSize, nThreads=0, extension=1) -- ntt_goldilocks.hpp:49-136.

    The twiddle tables the reference builds here (2^s sequential products) live in the library's
    per-context cache, built on first use; nThreads has no meaning on the GPU.
    """

    def __init__(self, maxDomainSize, nThreads=0, extension=1, ctx=None):
        self.s = 0
        self.extension = extension
        self._ctx = ctx
        if maxDomainSize == 0:
            return                                     # inert object (ntt_goldilocks.hpp:55-56)
        domainPow = int(maxDomainSize).bit_length() - 1
        if domainPow > 32:
            raise ValueError("Domain size too big for the curve")   # std::range_error, :94-97
        self.s = max(1, domainPow)

    @property
    def ctx(self):
        if self._ctx is None:
            self._ctx = default_context()
        return self._ctx

    def NTT(self, dst, src, size, ncols=1, buffer=None, nphase=NUM_PHASES, nblock=NUM_BLOCKS, inverse=False, extend=False):
        """ntt_goldilocks.cpp:166-232.  buffer/nphase/nblock are performance hints there; ignored here."""
        if ncols == 0 or size == 0:
            return
        if size & (size - 1):
            raise ValueError("size must be a power of two")         # assert, ntt_goldilocks.cpp:28
        self.ctx.ntt(dst, src, size, ncols, inverse, extend)

    def INTT(self, dst, src, size, ncols=1, buffer=None, nphase=NUM_PHASES, nblock=NUM_BLOCKS, extend=False):
        """ntt_goldilocks.cpp:307-324; dst None -> in place."""
        if ncols == 0 or size == 0:
            return
        self.NTT(src if dst is None else dst, src, size, ncols, buffer, nphase, nblock, True, extend)

    def extendPol(self, output, input, N_Extended, N, ncols, buffer=None, nphase=NUM_PHASES, nblock=NUM_BLOCKS):
        """ntt_goldilocks.cpp:326-352; output may alias input."""
        self.ctx.extend_pol(output, input, N_Extended, N, ncols)

    def reversePermutation(self, dst, src, size, offset_cols, ncols, ncols_all):
        """ntt_goldilocks.cpp:243-305."""
        self.ctx.reverse_permutation(dst, src, size, offset_cols, ncols, ncols_all, self.extension)

    def root(self, domainPow, idx):
        """ntt_goldilocks.hpp:169-172: W[domainPow]^idx."""
        return load().glp_root(domainPow, idx)

    def computeR(self, N):
        """ntt_goldilocks.hpp:153-165: the shift^i / N table is part of the library's cache; nothing to do."""
        return None


class PoseidonGoldilocks:
    """Static functions of PoseidonGoldilocks (poseidon_goldilocks.hpp:57-66) on the GPU."""

    @staticmethod
    def hash_full_result(state, input, ctx=None):
        (ctx or default_context()).hash_full_result(state, input, 1)

    @staticmethod
    def hash(state, input, ctx=None):
        tmp = np.zeros(12, dtype=np.uint64)
        (ctx or default_context()).hash_full_result(tmp, input, 1)
        state[:4] = tmp[:4]

    @staticmethod
    def linear_hash(output, input, size, ctx=None):
        (ctx or default_context()).linear_hash_rows(output, input, size, max(size, 1), 1)

    @staticmethod
    def merkletree(tree, input, num_cols, num_rows, nThreads=0, dim=1, ctx=None):
        if num_rows == 0:
            return
        (ctx or default_context()).merkletree(tree, input, num_cols, num_rows, dim)

    @staticmethod
    def merkletree_batch(tree, input, num_cols, num_rows, batch_size, nThreads=0, dim=1, ctx=None):
        """poseidon_goldilocks.hpp:44,82-88"""
        if num_rows == 0:
            return
        (ctx or default_context()).merkletree_batch(tree, input, num_cols, num_rows, batch_size, dim)

    merkletree_batch_avx = merkletree_batch_avx512 = merkletree_batch_seq = merkletree_batch
    merkletree_avx = merkletree
    merkletree_avx512 = merkletree
    merkletree_seq = merkletree


class MerkleTreeGL:
    """MerkleTreeGL(height, width, source) -- merkleTreeGL.hpp:33-43.

    `source` (height x width uint64, row-major) is aliased, or allocated when None; `nodes` is owned.
    After merkelize() both are host-visible numpy arrays, as FRI expects (fri/friProve.cpp:225,241).
    """

    def __init__(self, height=0, width=0, source=None, ctx=None, tree=None):
        self._ctx = ctx
        if tree is not None:                       # MerkleTreeGL(Goldilocks::Element *tree), merkleTreeGL.hpp:24-32
            tree = np.ascontiguousarray(tree, dtype=np.uint64)
            self.width, self.height = int(tree[0]), int(tree[1])
            self.source = tree[2:2 + self.height * self.width].reshape(self.height, self.width)
            self.nodes = tree[2 + self.height * self.width:]
            self.isSourceAllocated = self.isNodesAllocated = False
            return
        self.height, self.width = height, width
        self.isSourceAllocated = source is None
        self.source = np.zeros((height, width), dtype=np.uint64) if source is None else source
        self.nodes = np.zeros(self.getTreeNumElements(), dtype=np.uint64)
        self.isNodesAllocated = True

    @property
    def ctx(self):
        if self._ctx is None:
            self._ctx = default_context()
        return self._ctx

    def copySource(self, source):
        np.copyto(self.source.reshape(-1), np.asarray(source, dtype=np.uint64).reshape(-1))

    def getTreeNumElements(self):
        return self.height * HASH_SIZE + (self.height - 1) * HASH_SIZE

    def merkelize(self):
        """merkleTreeGL.cpp:37-44."""
        self.ctx.merkletree(self.nodes, self.source, self.width, self.height, 1)

    def getRoot(self, root=None):
        n = self.getTreeNumElements()
        r = self.nodes[n - HASH_SIZE:n].copy()
        if root is not None:
            root[:HASH_SIZE] = r
        return r

    def MerkleProofSize(self):
        return int(math.ceil(math.log10(self.height) / math.log10(2))) if self.height > 1 else 0

    def getGroupProof(self, proof, idx):
        """merkleTreeGL.cpp:12-35: row idx, then the sibling path."""
        self.ctx.merkle_group_proof(proof, self.source, self.nodes, self.width, self.height, idx)
        return proof


class Transcript:
    """Fiat-Shamir transcript of the STARK (src/starkpil/transcript/transcript.{hpp,cpp}): a Poseidon sponge that
    absorbs 8 words at a time (`pending`), keeps 4 words of `state`, and hands out the 12 outputs of the last
    permutation in order.  Same members and methods as the reference class; every permutation is one
    glp_hash_full_result call on the GPU (the host keeps only the cursors)."""

    STATE_SIZE, PENDING_SIZE, OUT_SIZE = 4, 8, 12          # transcript.hpp:8-10

    def __init__(self, ctx=None):
        self._ctx = ctx
        self.state = np.zeros(self.STATE_SIZE, dtype=np.uint64)
        self.pending = np.zeros(self.PENDING_SIZE, dtype=np.uint64)
        self.out = np.zeros(self.OUT_SIZE, dtype=np.uint64)
        self.pending_cursor = 0
        self.out_cursor = 0
        self.state_cursor = 0

    @property
    def ctx(self):
        return self._ctx or default_context()

    def _permute(self):
        """transcript.cpp:19-29 / :45-55: out = hash_full_result(pending || state); state = out[0:4]; pending cleared"""
        inputs = np.concatenate([self.pending, self.state])
        self.ctx.hash_full_result(self.out, inputs, 1)
        self.out_cursor = self.OUT_SIZE
        self.pending[:] = 0
        self.pending_cursor = 0
        self.state[:] = self.out[:self.STATE_SIZE]

    def _add1(self, value):
        self.pending[self.pending_cursor] = value
        self.pending_cursor += 1
        self.out_cursor = 0
        if self.pending_cursor == self.PENDING_SIZE:
            self._permute()

    def put(self, input, size=None):
        a = np.asarray(input, dtype=np.uint64).reshape(-1)
        for v in a[: a.size if size is None else size]:
            self._add1(v)

    def getFields1(self):
        if self.out_cursor == 0:
            self._permute()
        res = self.out[(self.OUT_SIZE - self.out_cursor) % self.OUT_SIZE]
        self.out_cursor -= 1
        return int(res)

    def getField(self, output=None):
        vals = [self.getFields1() for _ in range(3)]
        if output is not None:
            output[:3] = vals
        return vals

    def getPermutations(self, n, nBits):
        """transcript.cpp:58-91: n indices of nBits bits cut from the low 63 bits of successive field elements"""
        total_bits = n * nBits
        nfields = (total_bits - 1) // 63 + 1
        fields = [self.getFields1() for _ in range(nfields)]
        res, cur_field, cur_bit = [], 0, 0
        for _ in range(n):
            a = 0
            for j in range(nBits):
                if (fields[cur_field] >> cur_bit) & 1:
                    a += 1 << j
                cur_bit += 1
                if cur_bit == 63:
                    cur_bit, cur_field = 0, cur_field + 1
            res.append(a)
        return res


def build_const_tree(const_pols, nBits, nBitsExt, ctx=None):
    """bctree's buildConstTree for verificationHashType GL (tools/starkpil/bctree/build_const_tree.cpp:531-575):
    returns the *.consttree words [nPols, nExt, extended constant polynomials, Merkle nodes]; the constRoot of
    verkey.json is its last four words.  MerkleTreeGL(tree=...) maps it back (merkleTreeGL.hpp:24-32)."""
    ctx = ctx or default_context()
    a = np.ascontiguousarray(const_pols, dtype=np.uint64)
    n, n_ext = 1 << nBits, 1 << nBitsExt
    npols = a.size // n
    out = np.zeros(2 + npols * n_ext + tree_num_elements(n_ext), dtype=np.uint64)
    ctx.const_tree(out, a, n, n_ext, npols)
    return out
