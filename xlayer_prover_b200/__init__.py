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

from ._capi import (Context, GlpError, GlpLib, P, load, ptr_of, splitmix, synth_matrix, tree_num_elements,
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
