/* glprover.h -- C ABI of libglprover.so, the B200 (sm_100a) implementation of the zkEVM prover's
 * STARK-commitment hot path: Goldilocks NTT / INTT / extendPol + Poseidon-Goldilocks linear_hash / merkletree.
 *
 * The reference (okx/xlayer-prover) has no FFI layer on this path: Starks::genProof calls two C++ classes
 * directly.  The drop-in is therefore source level -- include/ntt_goldilocks.hpp and include/merkleTreeGL.hpp
 * keep the reference's class and method signatures and forward here.  Every entry point
 * below names the reference function it replaces (paths relative to the reference root).
 *
 * Conventions
 *  - Field elements are raw uint64_t, the same bytes as Goldilocks::Element{uint64_t fe}
 *    (src/goldilocks/src/goldilocks_base_field.hpp:24-27).  Inputs may be non-canonical (>= p);
 *    every output is canonical, i.e. already passed through Goldilocks::toU64
 *    (src/goldilocks/src/goldilocks_base_field_tools.hpp:146-163).
 *  - Matrices are row-major, element (row i, column c) at base[i*pitch + c], exactly the reference layout
 *    (src/goldilocks/src/ntt_goldilocks.cpp:95-109).
 *  - Every data pointer may be a device pointer or a host pointer (pageable or pinned); the library asks
 *    the driver which (cudaPointerGetAttributes) and stages host data through the GPU.  There is NO CPU
 *    implementation behind this ABI: without a usable CUDA device glp_init fails with GLP_E_CUDA.
 *  - Every function returns 0 or a negative GLP_E_* code; glp_last_error() gives the thread's last message.
 *  - A glp_ctx owns one device, one stream, its twiddle tables and scratch; it is not thread-safe, use one
 *    per host thread (the reference's NTT_Goldilocks objects are per-thread too, fri/friProve.cpp:93-105).
 */
#ifndef GLPROVER_H
#define GLPROVER_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GLP_OK 0
#define GLP_E_ARG (-1)      /* bad argument: NULL, non power-of-two size ...  (reference: assert, ntt_goldilocks.cpp:28) */
#define GLP_E_DOMAIN (-2)   /* domain larger than 2^32 (reference: std::range_error, ntt_goldilocks.hpp:94-97) */
#define GLP_E_CUDA (-3)     /* CUDA runtime failure or no device */
#define GLP_E_OOM (-4)      /* device or pinned-host allocation failed */

typedef struct glp_ctx glp_ctx;

/* ---- lifetime ---- */
int glp_init(int device, glp_ctx **ctx);              /* binds to CUDA device `device`, creates the stream */
void glp_destroy(glp_ctx *ctx);
const char *glp_last_error(void);
/* Launch on a caller-owned CUDA stream (cudaStream_t passed as void*); NULL restores the context's own. */
int glp_set_stream(glp_ctx *ctx, void *cuda_stream);
int glp_synchronize(glp_ctx *ctx);
/* kernels launched by this context so far (bench.py reports the difference as gpu_launches) */
uint64_t glp_launch_count(const glp_ctx *ctx);
/* leaf rows the hash kernel has had to redo with its exact arithmetic since the library was loaded (the fast path leaves a
 * 2^-33-per-multiplication wrap uncorrected and detects it); synchronises the context.  Diagnostics only. */
uint64_t glp_leaf_redo_rows(glp_ctx *ctx);
/* low-degree extensions this context has had to repeat with the exact NTT passes after the optimistic ones flagged a
 * left-out wrap (2^-64 per butterfly on honest data).  Diagnostics only. */
uint64_t glp_ntt_retry_count(const glp_ctx *ctx);
/* cap on internal scratch in bytes (0 = default: a quarter of free device memory, at most 16 GiB) */
int glp_set_scratch_limit(glp_ctx *ctx, uint64_t bytes);

/* frees the context's internal device buffers (staging, scratch, private result copies); they come back on demand */
int glp_release_scratch(glp_ctx *ctx);

/* ---- memory helpers (for callers that keep data resident on the device) ---- */
int glp_device_alloc(glp_ctx *ctx, uint64_t bytes, void **ptr);
int glp_device_free(glp_ctx *ctx, void *ptr);
int glp_host_alloc(glp_ctx *ctx, uint64_t bytes, void **ptr);     /* pinned */
int glp_host_free(glp_ctx *ctx, void *ptr);
/* pin / unpin memory the caller owns (e.g. the prover's polynomial buffer, src/prover/prover.cpp:101-115), so that the
 * host-pointer entry points below copy at the pinned rate */
int glp_host_register(glp_ctx *ctx, void *ptr, uint64_t bytes);
int glp_host_unregister(glp_ctx *ctx, void *ptr);
int glp_memcpy(glp_ctx *ctx, void *dst, const void *src, uint64_t bytes);   /* any direction, synchronous */
int glp_mem_info(glp_ctx *ctx, uint64_t *free_bytes, uint64_t *total_bytes);

/* ---- NTT_Goldilocks (src/goldilocks/src/ntt_goldilocks.{hpp,cpp}) ---- */
/* NTT_Goldilocks::NTT / INTT (ntt_goldilocks.cpp:166-232, :307-324).
 * dst[k,:] = sum_i src[i,:] w^(ik); inverse: dst[j,:] = 1/size * sum_k src[k,:] w^(-jk), times shift^j when
 * `extend` (the reference's INTT(..., extend=true), ntt_goldilocks.cpp:122-135).
 * size must be a power of two <= 2^32; size == 0 or ncols == 0 is a no-op; dst == NULL or dst == src works
 * in place.  The reference's buffer/nphase/nblock arguments are performance hints and have no equivalent. */
int glp_ntt(glp_ctx *ctx, uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, int inverse, int extend);
/* The same with rows >= live_rows of src READ AS ZERO (never loaded): what a transform through an NTT_Goldilocks object
 * built with extension > 1 computes, whose reversePermutation zero-fills every row whose source index is >= size/extension
 * (ntt_goldilocks.cpp:259-277).  live_rows >= size is glp_ntt. */
int glp_ntt_live(glp_ctx *ctx, uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, int inverse, int extend,
                 uint64_t live_rows);
/* NTT_Goldilocks::extendPol (ntt_goldilocks.cpp:326-352): out[i,c] = P_c(shift * w_ext^i), i < n_ext, where
 * P_c interpolates in[:,c] on the size-n subgroup.  out may alias in. */
int glp_extend_pol(glp_ctx *ctx, uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols);
/* Same on a column window: reads in[i*in_pitch + c], writes out[i*out_pitch + c], c < ncols.  This is the unit
 * of the unfused multi-GPU path (each rank extends its own column slice; north_star / SURVEY 8e).  Device pointers only.
 * STREAM ORDERED: unlike every other compute entry point this one does NOT synchronise -- the kernels are queued on the
 * context's stream (glp_set_stream) and the call returns; read `out` from that stream, or after glp_synchronize.  It always
 * runs the exact NTT passes (never the optimistic ones, which need a host check of their flag). */
int glp_extend_pol_pitched(glp_ctx *ctx, uint64_t *out, uint64_t out_pitch, const uint64_t *in, uint64_t in_pitch,
                           uint64_t n_ext, uint64_t n, uint64_t ncols);
/* The same extension with its OUTPUT row-sharded: row i of the result (i < n_ext) is written to
 * parts[i / R] + (i % R) * part_pitch + part_col0, R = n_ext / nparts (nparts a power of two).  This is the multi-GPU
 * exchange of SURVEY 8e fused into the transform: rank g passes the row slabs of all ranks (its own, and its peers'
 * mapped with glp_ipc_import), part_col0 = its first column, and the last NTT pass stores every tile straight into the
 * owner's reference-layout [R x part_pitch] slab over NVLink -- no all-to-all, no re-interleave pass.  A host `in` is
 * uploaded column block by column block under the transform of the previous block; a device `in` must not overlap any
 * slab (the transform may be issued twice, see glp_ntt_retry_count).  Either way the call returns when this rank's
 * stores have been issued and its stream is idle; the caller still fences the peers.  `parts` is a host array. */
int glp_extend_pol_scatter(glp_ctx *ctx, uint64_t *const *parts, uint32_t nparts, uint64_t part_pitch, uint64_t part_col0,
                           const uint64_t *in, uint64_t in_pitch, uint64_t n_ext, uint64_t n, uint64_t ncols);
/* CUDA IPC plumbing for the above (one process per GPU): export a glp_device_alloc'ed buffer as 64 opaque bytes,
 * map a peer's buffer into this process, unmap it. */
int glp_ipc_export(glp_ctx *ctx, void *devptr, unsigned char handle[64]);
int glp_ipc_import(glp_ctx *ctx, const unsigned char handle[64], void **devptr);
int glp_ipc_close(glp_ctx *ctx, void *devptr);
/* NTT_Goldilocks::reversePermutation (ntt_goldilocks.cpp:243-305): dst[i, 0:ncols] = src[BR(i), offset_cols : offset_cols+ncols]
 * (dst rows have ncols words, src rows ncols_all).  With extension > 1 rows whose source index is >= size/extension
 * are zero-filled.  dst == src permutes in place (requires ncols == ncols_all, offset_cols == 0, extension <= 1). */
int glp_reverse_permutation(glp_ctx *ctx, uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t offset_cols,
                            uint64_t ncols, uint64_t ncols_all, uint64_t extension);
/* NTT_Goldilocks::root(domainPow, idx) = W[domainPow]^idx (ntt_goldilocks.hpp:169-172), canonical. Host only. */
uint64_t glp_root(uint32_t domain_pow, uint64_t idx);

/* ---- PoseidonGoldilocks (src/goldilocks/src/poseidon_goldilocks.{hpp,cpp}) ---- */
/* hash_full_result (poseidon_goldilocks.cpp:166-228) on n independent 12-word states; out may alias in. */
int glp_hash_full_result(glp_ctx *ctx, uint64_t *out, const uint64_t *in, uint64_t n);
/* linear_hash (poseidon_goldilocks.cpp:229-265) of every row: out[i*4..+4) = linear_hash(in[i*pitch..+ncols)). */
int glp_linear_hash_rows(glp_ctx *ctx, uint64_t *out4, const uint64_t *in, uint64_t ncols, uint64_t pitch, uint64_t nrows);
/* The same sponge one column block at a time (poseidon_goldilocks.cpp:240-262 is a loop over 8-word blocks whose
 * only carried state is the 4-word capacity = the previous output): absorbs in[i*pitch..+ncols) into state4[i*4..+4).
 * first != 0 starts a row (capacity zero; a whole row of <= 4 words is copied, :234-239); every block but a row's last
 * must be a multiple of 8 columns.  After the last block state4 holds linear_hash of the whole row.  This is what
 * lets glp_commit hash column blocks while later blocks are still being uploaded and extended. */
int glp_linear_hash_absorb(glp_ctx *ctx, uint64_t *state4, const uint64_t *in, uint64_t ncols, uint64_t pitch, uint64_t nrows, int first);
/* merkletree / merkletree_avx / merkletree_avx512 (poseidon_goldilocks.cpp:266-301, :484-520):
 * tree = (2*nrows-1)*4 words, level-major, root last.  A row is ncols*dim words.  nrows must be a power of
 * two (the reference reads a neighbour's node otherwise; every caller passes powers of two). */
int glp_merkletree(glp_ctx *ctx, uint64_t *tree, const uint64_t *in, uint64_t ncols, uint64_t nrows, uint64_t dim);
/* merkletree_batch / _avx / _avx512 (poseidon_goldilocks.cpp:113-164, :302-352, :521-577): the leaf of a row is the
 * linear_hash of the concatenated linear_hashes of its column batches (batch_size columns each, the last one shorter);
 * the levels above are the same.  Not called by MerkleTreeGL / Starks; provided with its known answers. */
int glp_merkletree_batch(glp_ctx *ctx, uint64_t *tree, const uint64_t *in, uint64_t ncols, uint64_t nrows, uint64_t batch_size, uint64_t dim);
/* The level loop alone (poseidon_goldilocks.cpp:283-300): `level` holds n nodes (4 words each) followed by
 * room for the n-1 nodes above them.  Used for the top log2(G) levels of the multi-GPU tree. */
int glp_merkle_levels(glp_ctx *ctx, uint64_t *level, uint64_t n);
/* MerklehashGoldilocks::getTreeNumElements (merklehash_goldilocks.hpp:24-27) */
uint64_t glp_tree_num_elements(uint64_t nrows);
/* MerkleTreeGL::getGroupProof (src/starkpil/merkleTree/merkleTreeGL.cpp:12-35): row idx of source followed by
 * the sibling path; proof has width + 4*ceil(log2 height) words.  source / nodes may live on the device. */
int glp_merkle_group_proof(glp_ctx *ctx, uint64_t *proof, const uint64_t *source, const uint64_t *nodes,
                           uint64_t width, uint64_t height, uint64_t idx);
/* The same for a batch of indices, in one launch and one download: proofs[q * (width + 4*ceil(log2 height)) ...] =
 * getGroupProof(idx[q]), q < nidx.  This is the FRI query phase's access pattern (src/starkpil/fri/friProve.cpp:150-170,
 * 219-250: every tree is asked for every query index); with source and nodes on the device only the proofs cross the link.
 * idx is a host array; source and nodes are both device or both host memory; proofs may be either when they are device. */
int glp_merkle_group_proofs(glp_ctx *ctx, uint64_t *proofs, const uint64_t *source, const uint64_t *nodes, uint64_t width,
                            uint64_t height, const uint64_t *idx, uint64_t nidx);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  glp_merkle_group_proofs with the staged
 * query indices kept in a file-scope device array of GLP_GROUP_PROOFS_STATIC_MAX entries (32 KiB, held by every GPU that
 * loads the library until exit) instead of the context's growable scratch: it calls no cudaMalloc / cudaFree and no
 * allocating library routine.  proofs, source and nodes must be DEVICE memory (the caller owns inputs and outputs), idx is a
 * host array; nidx > GLP_GROUP_PROOFS_STATIC_MAX is rejected with GLP_E_ARG.  nidx == 0 performs only the per-device set-up
 * (the array's address is looked up once per device), so it can be done ahead of time.  All calls on one device share the
 * array, which keeps the previous call's indices (each call overwrites the entries it reads; nothing needs zeroing): calls
 * that overlap on the GPU -- two contexts or streams of one device -- race; issue them one after another.  Returns with
 * the context's stream idle. */
#define GLP_GROUP_PROOFS_STATIC_MAX 4096
int glp_merkle_group_proofs_static(glp_ctx *ctx, uint64_t *proofs, const uint64_t *source, const uint64_t *nodes, uint64_t width,
                                   uint64_t height, const uint64_t *idx, uint64_t nidx);

/* ---- the fused hot path: Starks::genProof step 1 (src/starkpil/starks.cpp:53-57) ----
 * out_ext = extendPol(in), tree = merkletree(out_ext), root = last 4 words of tree, with no host round trip
 * in between.  out_ext and tree may be NULL when the caller only needs the root (they are then kept in
 * device scratch).  With a host `in` the commit is a pipeline over column blocks (GLP_STREAM_COLS columns, default 160):
 * upload of block k+1, LDE + sponge absorption of block k and download of block k-1 overlap on three streams. */
int glp_commit(glp_ctx *ctx, uint64_t *out_ext, uint64_t *tree, uint64_t root[4], const uint64_t *in,
               uint64_t n, uint64_t n_ext, uint64_t ncols);

/* ---- several GPUs, one process (SURVEY 8e behind the C ABI) ----
 * The reference's prover is one C++ process (Starks::genProof, src/starkpil/starks.cpp:53-57,134-138,215-219); a glp_multi
 * drives G = 2^k devices from it.  glp_multi_init creates one glp_ctx per device and enables peer access between all of
 * them (NVLink/NVSwitch).  Every call below takes HOST pointers (pinned memory copies fastest) and returns when the
 * results are in them; work is split exactly as in the one-process-per-GPU path (xlayer_prover_b200/sharded.py):
 * columns of the trace for the LDE, rows of the extension for the tree, the exchange in between fused into the LDE's
 * last NTT pass as stores into the owning device's row slab, the top log2(G) tree levels on the first device.
 * Results are bit-identical for every G.  devs == NULL means devices 0..ndev-1; listing ONE device ndev times is allowed
 * (testing on a one-GPU box).  Shapes too small to split (fewer than 8 columns or 2 rows per device), and calls handed a
 * device pointer (device-resident data belongs to one device), run on the first device's context. */
typedef struct glp_multi glp_multi;
int glp_multi_init(int ndev, const int *devs, glp_multi **m);
void glp_multi_destroy(glp_multi *m);
int glp_multi_device_count(const glp_multi *m);
glp_ctx *glp_multi_ctx(glp_multi *m, int i);            /* the i-th device's context (owned by m) */
/* NTT_Goldilocks::extendPol (ntt_goldilocks.cpp:326-352), column-sharded: device g extends columns [c_g, c_{g+1}). */
int glp_multi_extend_pol(glp_multi *m, uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols);
/* PoseidonGoldilocks::merkletree (poseidon_goldilocks.cpp:266-301), row-sharded: device g hashes rows [g R, (g+1) R). */
int glp_multi_merkletree(glp_multi *m, uint64_t *tree, const uint64_t *in, uint64_t ncols, uint64_t nrows, uint64_t dim);
/* glp_commit on G devices: extendPol + merkletree with the extension exchanged over NVLink inside the LDE. */
int glp_multi_commit(glp_multi *m, uint64_t *out_ext, uint64_t *tree, uint64_t root[4], const uint64_t *in, uint64_t n,
                     uint64_t n_ext, uint64_t ncols);

/* One step of the above for callers that run one process per GPU (xlayer_prover_b200/sharded.py): copy the levels of this
 * rank's subtree (sub_rows leaves, level-major) into their slices of the full level-major tree of sub_rows * nparts leaves at
 * `tree` -- typically rank 0's buffer mapped with glp_ipc_import.  STREAM ORDERED: queued on the context's stream, not
 * synchronised; the caller fences the ranks before the top levels are hashed (glp_merkle_levels on the nparts subtree roots). */
int glp_place_subtree(glp_ctx *ctx, uint64_t *tree, const uint64_t *subtree, uint64_t sub_rows, uint32_t nparts, uint32_t part);

/* ---- the callers either side of the commit ---- */
/* Step 4 of Starks::genProof, the quotient commit (src/starkpil/starks.cpp:261-291):
 *   qq1 = INTT_{n_ext}(q)  (q: n_ext x qdim);  qq2[k][p*qdim+d] = qq1[p*n+k][d] * shift^(-n*p)  (k < n, p < qdeg);
 *   cm4 = NTT_{n_ext}(qq2 zero-padded)  (n_ext x qdim*qdeg);  tree = merkletree(cm4).
 * cm4 / tree / root may be NULL (not wanted); host or device pointers. */
int glp_quotient_commit(glp_ctx *ctx, uint64_t *cm4, uint64_t *tree, uint64_t root[4], const uint64_t *q, uint64_t n,
                        uint64_t n_ext, uint64_t qdim, uint64_t qdeg);
/* The constant tree as bctree writes it and MerkleTreeGL(Goldilocks::Element *tree) maps it
 * (tools/starkpil/bctree/build_const_tree.cpp:555-575; src/starkpil/merkleTree/merkleTreeGL.hpp:24-32):
 * const_tree = [npols, n_ext, extendPol(const_pols) (n_ext x npols), Merkle nodes]; the root is its last 4 words. */
uint64_t glp_const_tree_num_elements(uint64_t n_ext, uint64_t npols);
int glp_const_tree(glp_ctx *ctx, uint64_t *const_tree, const uint64_t *const_pols, uint64_t n, uint64_t n_ext, uint64_t npols);

/* ---- FRI layers on the device ---- */
/* One folding step of FRIProve::prove (src/starkpil/fri/friProve.cpp:22-108, si > 0): pol holds 2^pol_bits elements of
 * the cubic extension (3 words each); out[g] = evalPol(polMulAxi(INTT_{nX}(pol[i*2^step_bits + g]), shift_inv * w^-g), special_x),
 * g < 2^step_bits, nX = 2^(pol_bits - step_bits) <= 64. */
int glp_fri_fold(glp_ctx *ctx, uint64_t *out, const uint64_t *pol, uint32_t pol_bits, uint32_t step_bits, const uint64_t special_x[3],
                 uint64_t shift_inv);
/* The layer's tree (friProve.cpp:110-123, getTransposed :252-270): aux = pol regrouped into 2^next_bits rows,
 * tree = MerkleTreeGL(2^next_bits, 3 * 2^(pol_bits - next_bits), aux).  aux / tree / root may be NULL. */
int glp_fri_group_tree(glp_ctx *ctx, uint64_t *aux, uint64_t *tree, uint64_t root[4], const uint64_t *pol, uint32_t pol_bits,
                       uint32_t next_bits);

/* ---- synthetic data and checks used by bench.py / tests ---- */
/* dst[i*ncols + c] = splitmix64(seed ^ ((row0+i)*ncols_total + col0 + c)) mod p, generated on the device */
int glp_synth_fill(glp_ctx *ctx, uint64_t *dst, uint64_t nrows, uint64_t ncols, uint64_t row0, uint64_t col0,
                   uint64_t ncols_total, uint64_t seed);
/* dst[j*ncols + i] = (i+1) * F_j mod p, F_0 = F_1 = 1, F_j = F_(j-2) + F_(j-1): the Fibonacci matrix of the reference's Merkle
 * benchmark and tests (src/goldilocks/benchs/bench.cpp:410-427), generated on the device */
int glp_fib_fill(glp_ctx *ctx, uint64_t *dst, uint64_t nrows, uint64_t ncols);
/* out[0] = sum of all words mod p, out[1] = sum of word_i * (i+1) mod p  (order-sensitive checksum) */
int glp_checksum(glp_ctx *ctx, const uint64_t *src, uint64_t n, uint64_t out[2]);
/* exercises every device field primitive (PTX paths) on n pseudo-random + edge operands and returns the
 * results for the host to compare against the oracle:
 * out[8*i..] = a+b, a-b, a*b, a*2^(b%96), canon(a), a^2, b^7, add_c(a, canon(b)) */
int glp_field_selftest(glp_ctx *ctx, const uint64_t *a, const uint64_t *b, uint64_t n, uint64_t *out);

/* ---- measurement hooks (bench.py) ---- */
/* When on, the NTT and Merkle stages are bracketed by CUDA events on the context's stream. */
int glp_set_profiling(glp_ctx *ctx, int on);
/* Synchronises, then returns and clears the accumulated stage times in milliseconds and launch counts:
 * ms[0] inverse NTT (+coset scale), ms[1] forward NTT, ms[2] leaf hash (linear_hash rows), ms[3] tree levels;
 * count[i] = number of bracketed intervals summed into ms[i]. */
int glp_get_profile(glp_ctx *ctx, float ms[4], uint32_t count[4]);
/* Dependent-free integer micro-benchmarks on this device: ops_per_s[0] = IMAD (32-bit multiply-add, the IMAD.WIDE pipe) lane-ops/s,
 * [1] = 32-bit add lane-ops/s (two adds fused as IADD3 count as one), [2] = both interleaved (sum of lane-ops/s). */
int glp_measure_int_peak(glp_ctx *ctx, double ops_per_s[3]);
/* The same three plus [3] = IMAD.WIDE (32x32 -> 64 with a 64-bit accumulate) and [4] = DFMA lane-ops/s: the two instruction
 * classes the algorithmic floor of one Poseidon permutation is counted in (1122 field multiplies x 4 IMAD.WIDE + 7 MDS layers
 * x 288 DFMA; bench.py `roofline`). */
int glp_measure_pipe_peaks(glp_ctx *ctx, double ops_per_s[5]);

#ifdef __cplusplus
}
#endif
#endif
