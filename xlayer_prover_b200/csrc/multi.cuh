// Multi-GPU commit inside ONE process, behind the C ABI (glprover.h: glp_multi_*): what xlayer_prover_b200/sharded.py does
// with one process per GPU and torch.distributed, for callers that are a single C++ process -- the reference's prover is
// (Starks::genProof, src/starkpil/starks.cpp:53-57,134-138,215-219 calls ntt.extendPol and treesGL[i]->merkelize() from one
// thread).  SURVEY.md 8e:
//
//   * device g owns the column slice [c_g, c_{g+1}) of the trace (boundaries are multiples of 8, so every 8-word sponge block
//     of the leaf hash has one owner) and the row slab [g R, (g+1) R), R = N_ext / G, of the extension;
//   * the exchange is FUSED into the LDE: every device enables peer access to the others (cudaDeviceEnablePeerAccess: with
//     unified addressing a peer's cudaMalloc'ed slab is a plain pointer, no IPC handles), and the last NTT pass stores each
//     32-column x 256-row tile straight into the owner's slab at column offset c_g over NVLink (glp_extend_pol_scatter);
//   * device g hashes its R leaf rows, builds its subtree, and writes the subtree's levels straight into their slices of the
//     reference's level-major `nodes` array (poseidon_goldilocks.cpp:283-300) -- the caller's host buffer, or device 0's;
//   * device 0 hashes the top log2(G) levels from the G subtree roots.
//
// One host thread per device drives that device's context (every glp_ctx is single-threaded; they share nothing but the
// slabs), with host barriers where the per-process version has its one-word all-reduces.  Nothing here spins on the device.
#pragma once
#include <condition_variable>
#include <mutex>
#include <thread>

struct glp_multi {
    std::vector<glp_ctx *> ctx;
    std::vector<int> dev;
    std::vector<DevBuf> slab, sub;           // per device: its [R x C] row slab of the extension, its subtree
    DevBuf top;                              // device 0: the G subtree roots and the levels above them
    bool same_device = false;                // every entry names the same device (tests on a one-GPU box / the emulator)
};

namespace {
struct HostBarrier {
    std::mutex m;
    std::condition_variable cv;
    unsigned n, waiting = 0, gen = 0;
    explicit HostBarrier(unsigned n_) : n(n_) {}
    void wait()
    {
        std::unique_lock<std::mutex> lk(m);
        const unsigned g = gen;
        if (++waiting == n) { waiting = 0; gen++; cv.notify_all(); return; }
        cv.wait(lk, [&] { return gen != g; });
    }
};

// c_0..c_G, every boundary but the last a multiple of 8 (sharded.py column_bounds)
std::vector<uint64_t> multi_column_bounds(uint64_t ncols, unsigned G)
{
    const uint64_t nblk = (ncols + 7) / 8;
    std::vector<uint64_t> b(G + 1);
    for (unsigned g = 0; g < G; g++) { const uint64_t c = 8 * ((g * nblk) / G); b[g] = c < ncols ? c : ncols; }
    b[G] = ncols;
    return b;
}

// run fn(g) on one host thread per device; the first failure (lowest g) is returned and becomes the caller's last error
template <class F>
int multi_run(glp_multi *m, F &&fn)
{
    const unsigned G = (unsigned)m->ctx.size();
    std::vector<int> rc(G, GLP_OK);
    std::vector<std::string> err(G);
    if (G == 1) {
        const int r = fn(0u);
        return r;
    }
    std::vector<std::thread> th;
    for (unsigned g = 0; g < G; g++)
        th.emplace_back([&, g]() {
            rc[g] = fn(g);
            if (rc[g] != GLP_OK) err[g] = g_err;           // g_err is thread_local
        });
    for (auto &t : th) t.join();
    for (unsigned g = 0; g < G; g++)
        if (rc[g] != GLP_OK) return fail(rc[g], "device " + std::to_string(m->dev[g]) + ": " + err[g]);
    return GLP_OK;
}

// Copy the levels of a subtree of R leaves (device, level-major) into their slices of the full level-major tree of
// R * G leaves at `tree` (host or device; UVA), subtree index g.  The last level copied is the one with G nodes.
int multi_place_subtree(glp_ctx *ctx, uint64_t *tree, const gl_t *sub, uint64_t R, unsigned G, unsigned g, cudaStream_t st)
{
    uint64_t off_full = 0, off_sub = 0, n_full = R * G, n_sub = R;
    for (;;) {
        CU(cudaMemcpyAsync(tree + off_full + 4 * g * n_sub, sub + off_sub, 4 * n_sub * sizeof(gl_t), cudaMemcpyDefault, st));
        if (n_sub == 1) break;
        off_full += 4 * n_full;
        off_sub += 4 * n_sub;
        n_full >>= 1;
        n_sub >>= 1;
    }
    (void)ctx;
    return GLP_OK;
}
uint64_t multi_top_offset(uint64_t R, unsigned G)           // word offset of the level with G nodes in the full tree
{
    uint64_t off = 0, n = R * G;
    while (n > G) { off += 4 * n; n >>= 1; }
    return off;
}
}  // namespace

// The placement step on its own, for the one-process-per-GPU path (sharded.py): rank `part` of `nparts` copies the levels of its
// subtree (sub_rows leaves) into their slices of the full tree, which may be rank 0's buffer mapped over CUDA IPC.  Stream
// ordered: the copies are queued on the context's stream and the call returns; the caller fences the ranks.
extern "C" int glp_place_subtree(glp_ctx *ctx, uint64_t *tree, const uint64_t *subtree, uint64_t sub_rows, uint32_t nparts, uint32_t part)
{
    if (!ctx || !tree || !subtree) return fail(GLP_E_ARG, "glp_place_subtree: NULL argument");
    if (!is_pow2(sub_rows) || !is_pow2(nparts) || part >= nparts) return fail(GLP_E_ARG, "glp_place_subtree: bad shape");
    CU(cudaSetDevice(ctx->device));
    return multi_place_subtree(ctx, tree, (const gl_t *)subtree, sub_rows, nparts, part, ctx->stream);
}

extern "C" int glp_multi_init(int ndev, const int *devs, glp_multi **out)
{
    if (!out) return fail(GLP_E_ARG, "glp_multi_init: NULL argument");
    *out = nullptr;
    if (ndev < 1 || ndev > 64 || !is_pow2((uint64_t)ndev)) return fail(GLP_E_ARG, "glp_multi_init: the device count must be a power of two (1..64)");
    int avail = 0;
    cudaError_t e = cudaGetDeviceCount(&avail);
    if (e != cudaSuccess || avail <= 0) {
        cudaGetLastError();
        return fail(GLP_E_CUDA, std::string("glp_multi_init: no usable CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path");
    }
    glp_multi *m = new glp_multi();
    m->same_device = true;
    for (int i = 0; i < ndev; i++) {
        const int d = devs ? devs[i] : i;
        if (d < 0 || d >= avail) { delete m; return fail(GLP_E_ARG, "glp_multi_init: device index out of range"); }
        m->dev.push_back(d);
        if (d != m->dev[0]) m->same_device = false;
    }
    // the same device may be listed more than once only when it is the ONLY device listed (one-GPU boxes, the emulator)
    if (!m->same_device)
        for (int i = 0; i < ndev; i++)
            for (int j = 0; j < i; j++)
                if (m->dev[i] == m->dev[j]) { delete m; return fail(GLP_E_ARG, "glp_multi_init: a device is listed twice"); }
    for (int i = 0; i < ndev; i++) {
        glp_ctx *c = nullptr;
        const int rc = glp_init(m->dev[i], &c);
        if (rc != GLP_OK) { for (glp_ctx *x : m->ctx) glp_destroy(x); delete m; return rc; }
        m->ctx.push_back(c);
    }
    if (!m->same_device) {
        for (int i = 0; i < ndev; i++) {
            cudaSetDevice(m->dev[i]);
            for (int j = 0; j < ndev; j++) {
                if (i == j) continue;
                int can = 0;
                cudaDeviceCanAccessPeer(&can, m->dev[i], m->dev[j]);
                if (!can) {
                    for (glp_ctx *x : m->ctx) glp_destroy(x);
                    delete m;
                    return fail(GLP_E_CUDA, "glp_multi_init: devices " + std::to_string(m->dev[i]) + " and " + std::to_string(m->dev[j]) + " cannot access each other's memory");
                }
                const cudaError_t pe = cudaDeviceEnablePeerAccess(m->dev[j], 0);
                if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) {
                    for (glp_ctx *x : m->ctx) glp_destroy(x);
                    delete m;
                    return fail(GLP_E_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(pe));
                }
                cudaGetLastError();
            }
        }
    }
    m->slab.resize(ndev);
    m->sub.resize(ndev);
    *out = m;
    return GLP_OK;
}

extern "C" void glp_multi_destroy(glp_multi *m)
{
    if (!m) return;
    for (size_t g = 0; g < m->ctx.size(); g++) {
        cudaSetDevice(m->dev[g]);
        cudaStreamSynchronize(m->ctx[g]->stream);
    }
    for (size_t g = 0; g < m->ctx.size(); g++) {
        cudaSetDevice(m->dev[g]);
        release(m->slab[g]);
        release(m->sub[g]);
        if (g == 0) release(m->top);
        glp_destroy(m->ctx[g]);
    }
    delete m;
}
extern "C" int glp_multi_device_count(const glp_multi *m) { return m ? (int)m->ctx.size() : 0; }
extern "C" glp_ctx *glp_multi_ctx(glp_multi *m, int i) { return (m && i >= 0 && i < (int)m->ctx.size()) ? m->ctx[i] : nullptr; }

// NTT_Goldilocks::extendPol on G devices: columns are independent polynomials, so device g extends the column window
// [c_g, c_{g+1}) of the host matrix (upload, LDE and download of column blocks overlapped, commit_streamed) -- no exchange.
extern "C" int glp_multi_extend_pol(glp_multi *m, uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols)
{
    if (!m) return fail(GLP_E_ARG, "glp_multi_extend_pol: NULL context");
    if (n == 0 || ncols == 0) return GLP_OK;
    GLP(check_extend_args(out, in, n_ext, n));
    const unsigned G = (unsigned)m->ctx.size();
    // device-resident data belongs to one device: its context does the work
    if (G == 1 || ncols < 8 * (uint64_t)G || on_device(in) || on_device(out)) return glp_extend_pol(m->ctx[0], out, in, n_ext, n, ncols);
    const std::vector<uint64_t> b = multi_column_bounds(ncols, G);
    return multi_run(m, [&](unsigned g) -> int {
        glp_ctx *ctx = m->ctx[g];
        const uint64_t w = b[g + 1] - b[g];
        if (!w) return GLP_OK;
        CU(cudaSetDevice(ctx->device));
        return commit_streamed(ctx, out + b[g], false, nullptr, false, nullptr, nullptr, in + b[g], n, n_ext, w, ncols, ncols);
    });
}

// merkletree on G devices: rows are independent leaves, so device g uploads and hashes the row slab [g R, (g+1) R) of the
// host source, builds the subtree above it and writes its levels into their slices of `tree`; device 0 finishes the top.
static int multi_tree_from_slabs(glp_multi *m, uint64_t *tree, uint64_t root[4], uint64_t R, HostBarrier &bar, unsigned g,
                                 const gl_t *sub)
{
    const unsigned G = (unsigned)m->ctx.size();
    glp_ctx *ctx = m->ctx[g];
    const uint64_t top_off = multi_top_offset(R, G);
    const size_t top_words = (size_t)(2 * G - 1) * 4;
    int rc = GLP_OK;
    // every device hands its subtree root to device 0 (a 32-byte peer copy), then places its levels
    rc = [&]() -> int {
        CU(cudaMemcpyAsync((gl_t *)m->top.p + 4 * g, sub + glp_tree_num_elements(R) - 4, 4 * sizeof(gl_t), cudaMemcpyDefault, ctx->stream));
        if (tree) GLP(multi_place_subtree(ctx, tree, sub, R, G, g, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        return GLP_OK;
    }();
    bar.wait();                                                    // all G roots are on device 0
    if (rc != GLP_OK) return rc;
    if (g == 0) {
        CU(cudaSetDevice(ctx->device));
        {
            ProfScope ps(ctx, 3);
            GLP(dev_tree_levels(ctx, (gl_t *)m->top.p, G));
        }
        if (tree) CU(cudaMemcpyAsync(tree + top_off + 4 * G, (gl_t *)m->top.p + 4 * G, (top_words - 4 * G) * sizeof(gl_t), cudaMemcpyDefault, ctx->stream));
        if (root) CU(cudaMemcpyAsync(root, (gl_t *)m->top.p + top_words - 4, 4 * sizeof(gl_t), cudaMemcpyDefault, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    return GLP_OK;
}

extern "C" int glp_multi_merkletree(glp_multi *m, uint64_t *tree, const uint64_t *in, uint64_t ncols, uint64_t nrows, uint64_t dim)
{
    if (!m) return fail(GLP_E_ARG, "glp_multi_merkletree: NULL context");
    if (nrows == 0) return GLP_OK;
    const uint64_t width = ncols * dim;
    if (!tree || (!in && width)) return fail(GLP_E_ARG, "glp_multi_merkletree: NULL pointer");
    if (!is_pow2(nrows)) return fail(GLP_E_ARG, "glp_multi_merkletree: nrows must be a power of two");
    const unsigned G = (unsigned)m->ctx.size();
    if (G == 1 || nrows < 2 * (uint64_t)G || width <= POS_CAP || on_device(in) || on_device(tree)) return glp_merkletree(m->ctx[0], tree, in, ncols, nrows, dim);
    const uint64_t R = nrows / G;
    {
        glp_ctx *c0 = m->ctx[0];
        CU(cudaSetDevice(c0->device));
        GLP(ensure(c0, m->top, (size_t)(2 * G - 1) * 4 * sizeof(gl_t)));
    }
    HostBarrier bar(G);
    return multi_run(m, [&](unsigned g) -> int {
        glp_ctx *ctx = m->ctx[g];
        int rc = [&]() -> int {
            CU(cudaSetDevice(ctx->device));
            GLP(ensure(ctx, m->sub[g], glp_tree_num_elements(R) * sizeof(gl_t)));
            gl_t *sub = (gl_t *)m->sub[g].p;
            const uint64_t *slab = in + (size_t)g * R * width;
            const uint64_t stream_min = ctx->merkle_stream_min ? ctx->merkle_stream_min : GLP_MERKLE_STREAM_MIN_BYTES;
            if (R >= 2048 && R * width * sizeof(gl_t) >= stream_min) return merkletree_streamed(ctx, sub, slab, width, R);
            const gl_t *din = nullptr;
            GLP(stage_in(ctx, ctx->tmp_in, slab, (size_t)R * width, &din));
            return dev_merkletree(ctx, sub, din, width, width, R);
        }();
        if (rc != GLP_OK) { bar.wait(); return rc; }                 // keep the others' barrier count right
        return multi_tree_from_slabs(m, tree, nullptr, R, bar, g, (const gl_t *)m->sub[g].p);
    });
}

// The fused commit (glp_commit) on G devices, host buffers: see the header of this file.  out_ext / tree / root may be NULL.
extern "C" int glp_multi_commit(glp_multi *m, uint64_t *out_ext, uint64_t *tree, uint64_t root[4], const uint64_t *in, uint64_t n,
                                uint64_t n_ext, uint64_t ncols)
{
    if (!m) return fail(GLP_E_ARG, "glp_multi_commit: NULL context");
    if (n == 0 || ncols == 0) return GLP_OK;
    if (!in) return fail(GLP_E_ARG, "glp_multi_commit: in is NULL");
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GLP_E_ARG, "glp_multi_commit: sizes must be powers of two with n_ext >= n");
    if (n_ext > (1ull << 32)) return fail(GLP_E_DOMAIN, "Domain size too big for the curve");
    const unsigned G = (unsigned)m->ctx.size();
    if (G == 1 || n_ext < 2 * (uint64_t)G || ncols < 8 * (uint64_t)G || on_device(in) || on_device(out_ext) || on_device(tree))
        return glp_commit(m->ctx[0], out_ext, tree, root, in, n, n_ext, ncols);
    const bool want_tree = tree != nullptr || root != nullptr;
    const uint64_t R = n_ext / G;
    const std::vector<uint64_t> b = multi_column_bounds(ncols, G);
    // slabs first (single-threaded: the pointers must all be known before any device starts storing)
    std::vector<uint64_t *> parts(G);
    for (unsigned g = 0; g < G; g++) {
        glp_ctx *ctx = m->ctx[g];
        CU(cudaSetDevice(ctx->device));
        GLP(ensure(ctx, m->slab[g], (size_t)R * ncols * sizeof(gl_t)));
        if (want_tree) GLP(ensure(ctx, m->sub[g], glp_tree_num_elements(R) * sizeof(gl_t)));
        if (g == 0 && want_tree) GLP(ensure(ctx, m->top, (size_t)(2 * G - 1) * 4 * sizeof(gl_t)));
        parts[g] = (uint64_t *)m->slab[g].p;
    }
    HostBarrier bar(G);
    return multi_run(m, [&](unsigned g) -> int {
        glp_ctx *ctx = m->ctx[g];
        gl_t *slab = (gl_t *)m->slab[g].p, *sub = (gl_t *)m->sub[g].p;
        const uint64_t w = b[g + 1] - b[g];
        // 1. my columns: upload block by block, extend, the last NTT pass scatters the rows to their owners' slabs
        int rc = GLP_OK;
        if (w) rc = glp_extend_pol_scatter(ctx, parts.data(), G, ncols, b[g], in + b[g], ncols, n_ext, n, w);
        bar.wait();                                                // every device's stores have landed (each synchronised its stream)
        if (rc != GLP_OK) { if (want_tree) bar.wait(); return rc; }
        // 2. my rows: download the slab (contiguous rows of out_ext) on a copy stream under the leaf hash and the subtree
        rc = [&]() -> int {
            CU(cudaSetDevice(ctx->device));
            if (out_ext) CU(cudaMemcpyAsync(out_ext + (size_t)g * R * ncols, slab, (size_t)R * ncols * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->copy_stream[1]));
            if (want_tree) GLP(dev_merkletree(ctx, sub, slab, ncols, ncols, R));
            return GLP_OK;
        }();
        if (want_tree) {
            if (rc != GLP_OK) { bar.wait(); }
            else rc = multi_tree_from_slabs(m, tree, root, R, bar, g, sub);
        }
        const cudaError_t e1 = cudaStreamSynchronize(ctx->copy_stream[1]);
        if (rc == GLP_OK) CU(e1);
        return rc;
    });
}
