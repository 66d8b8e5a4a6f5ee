// C ABI of libglprover.so (include/glprover.h): context, twiddle-table cache, pass planning, host<->device
// staging, and the kernel launches for the NTT/LDE and Poseidon/Merkle paths.  No CPU implementation lives
// here: every entry point ends in kernels from ntt_kernels.cuh / poseidon_kernels.cuh.
#include "cuda_compat.h"
#include "../../include/glprover.h"
#include "poseidon_kernels.cuh"
#include "ntt_kernels.cuh"
#include "aux_kernels.cuh"
#include "fri_kernels.cuh"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <map>
#include <mutex>
#include <string>
#include <vector>

// ------------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }
#define CU(call)                                                                                           \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) return fail(GLP_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)
#define GLP(call)                \
    do {                         \
        int rc_ = (call);        \
        if (rc_ != GLP_OK) return rc_; \
    } while (0)

extern "C" const char *glp_last_error(void) { return g_err.c_str(); }

// ------------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------------
struct DevBuf {
    void *p = nullptr;
    size_t bytes = 0;
};
struct TwoLevel {
    gl_t *lo = nullptr, *hi = nullptr;
};

struct glp_ctx {
    int device = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    cudaStream_t copy_stream[3] = {nullptr, nullptr, nullptr};   // [0] uploads, [1] downloads, [2] second download stream (GLP_D2H_SPLIT)
    uint32_t d2h_split = 2;                  // the streamed commit downloads each column block as this many row ranges on separate streams
    cudaStream_t lde_stream = nullptr;       // high priority: the LDE of column block k+1 under the leaf hash of block k (commit_overlapped)
    // co-resident NTT passes (k_ntt_pass_co): > 0 while the launches of an LDE should use the small resident grid
    uint32_t co_active = 0;
    uint32_t co_threads = 256, co_ctas_per_sm = 1;   // GLP_CO_THREADS, GLP_CO_CTAS
    uint32_t co_cols = 96, co_first_cols = 32;       // GLP_CO_COLS, GLP_CO_FIRST: column-block widths of the overlapped commit
    int co_mode = 0;                         // GLP_CO: 0 = stages back to back (default: measured fastest, profiles/r01_overlap_experiment.txt), 1 = LDE(k+1) under hash(k) with the co-resident kernels, 2 = same with the full-size kernels
    int sm_count = 148;
    uint64_t launches = 0;
    uint64_t scratch_limit = 0;
    uint64_t stream_cols = 160;              // column-block width of the streamed (host-pointer) commit; GLP_STREAM_COLS
    uint32_t stream_grow = 200;              // graded first blocks of the streamed commit, percent per step (0: one short block, then equal); GLP_STREAM_GROW
    std::vector<uint64_t> stream_plan;       // explicit block widths in columns; GLP_STREAM_PLAN
    uint32_t max_logq = NTT_MAX_LOGQ;        // tile size cap; GLP_NTT_MAX_LOGQ lowers it so tests reach the multi-pass paths at small sizes
    bool no_pass_tables = false;             // GLP_NTT_NO_PASS_TABLES=1: two-level twiddle/scale tables only (the path sizes > 2^25 take)
    bool trace = false;                      // GLP_TRACE=1: timeline of the streamed commit on stderr
    uint64_t merkle_stream_min = 0;          // GLP_MERKLE_STREAM_MIN: bytes from which a host-resident Merkle source is uploaded block by block (tests lower it)
    bool stage_all = true;                   // GLP_STAGE_ALL=0: two alternating upload buffers even when the whole input fits
    bool no_split = false;                   // GLP_NTT_NO_SPLIT=1: the remainder column group of a wide matrix stays on the lane = column mapping
    bool no_narrow = false;                  // GLP_NTT_NO_NARROW=1: lane = column even for matrices of <= 16 columns (comparison runs)
    // optimistic NTT passes (k_ntt_pass_opt): the device-resident LDE runs them and is repeated with the exact passes when
    // the flag they raise is set (dev_extend_pol_checked)
    bool opt_enabled = true;                 // GLP_NTT_OPT=0 turns them off
    bool opt_active = false;                 // true while dev_extend_pol_checked issues its first attempt
    bool opt_force_retry = false;            // GLP_NTT_FORCE_RETRY=1 (tests): behave as if the flag had been raised
    uint32_t *opt_flag_dev = nullptr;        // device word the kernels write
    uint32_t *opt_flag_host = nullptr;       // pinned copy
    uint64_t opt_retries = 0;                // transforms repeated with the exact passes (glp_ntt_retry_count)
    bool no_fused_lde = false;               // GLP_NTT_NO_FUSED_LDE=1: extendPol as two natural-order transforms (the fallback path)
    std::map<uint32_t, TwoLevel> tw_tabs;    // w_N^e, by logN
    std::map<uint32_t, TwoLevel> sc_tabs;    // shift^j / N, by logN
    std::map<uint64_t, gl_t *> pass_tabs;    // per-pass twiddle rows, by (logL, logQ, inverse)
    std::map<uint64_t, gl_t *> scale_tabs;   // shift^j / N in DIF storage order, by (logN, plan)
    gl_t *zeta = nullptr;                    // w_256^i
    DevBuf s1, s2;                           // NTT scratch
    DevBuf ext, nodes, tmp_in, tmp_out;      // commit / staging
    DevBuf partial;                          // checksum partials
    DevBuf work, parts_tab;                  // row-sharded LDE: dense block before its last pass, table of slab bases
    std::vector<void *> tables;              // everything to free
    bool profiling = false;
    struct ProfSpan { int kind; cudaEvent_t a, b; };
    std::vector<ProfSpan> spans;
};

// bracket a stage with events when profiling is on
struct ProfScope {
    glp_ctx *ctx;
    cudaEvent_t a = nullptr, b = nullptr;
    int kind;
    ProfScope(glp_ctx *c, int k) : ctx(c), kind(k)
    {
        if (!ctx->profiling) return;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        cudaEventRecord(a, ctx->stream);
    }
    ~ProfScope()
    {
        if (!a) return;
        cudaEventRecord(b, ctx->stream);
        ctx->spans.push_back({kind, a, b});
    }
};

// Events of a multi-stream pipeline (streamed commit / Merkle tree / scatter).  Whatever way the function is left -- the CU()
// and GLP() macros return early on an error -- the copy streams are drained before the events go away and before a later
// ensure() may free the staging buffers they still have work queued on.
struct PipelineEvents {
    glp_ctx *ctx;
    std::vector<cudaEvent_t> evs;
    explicit PipelineEvents(glp_ctx *c) : ctx(c) {}
    int make(cudaEvent_t *e, unsigned flags)
    {
        CU(cudaEventCreateWithFlags(e, flags));
        evs.push_back(*e);
        return GLP_OK;
    }
    ~PipelineEvents()
    {
        cudaStreamSynchronize(ctx->copy_stream[0]);
        cudaStreamSynchronize(ctx->copy_stream[1]);
        cudaStreamSynchronize(ctx->copy_stream[2]);
        cudaStreamSynchronize(ctx->stream);
        for (cudaEvent_t e : evs) cudaEventDestroy(e);
        cudaGetLastError();
    }
};

static int ensure(glp_ctx *ctx, DevBuf &b, size_t bytes)
{
    if (b.bytes >= bytes) return GLP_OK;
    if (b.p) { cudaFree(b.p); b.p = nullptr; b.bytes = 0; }
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        char m[160];
        snprintf(m, sizeof m, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
        return fail(GLP_E_OOM, m);
    }
    (void)ctx;
    b.p = p;
    b.bytes = bytes;
    return GLP_OK;
}
static void release(DevBuf &b)
{
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.bytes = 0;
}

#define LAUNCH(ctx, kernel, grid, block, smem, ...)                           \
    do {                                                                      \
        GLP_LAUNCH(kernel, grid, block, smem, (ctx)->stream, __VA_ARGS__);    \
        (ctx)->launches++;                                                    \
    } while (0)
#define LAUNCH_FLAT(ctx, kernel, grid, block, ...)                            \
    do {                                                                      \
        GLP_LAUNCH_FLAT(kernel, grid, block, (ctx)->stream, __VA_ARGS__);     \
        (ctx)->launches++;                                                    \
    } while (0)

extern "C" int glp_init(int device, glp_ctx **out)
{
    if (!out) return fail(GLP_E_ARG, "glp_init: ctx is NULL");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        return fail(GLP_E_CUDA, std::string("glp_init: no usable CUDA device (") + cudaGetErrorString(e) +
                                    "); this library has no CPU path");
    }
    if (device < 0 || device >= ndev) return fail(GLP_E_ARG, "glp_init: device index out of range");
    CU(cudaSetDevice(device));
    glp_ctx *ctx = new glp_ctx();
    ctx->device = device;
    CU(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&ctx->copy_stream[0], cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&ctx->copy_stream[1], cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&ctx->copy_stream[2], cudaStreamNonBlocking));
    ctx->stream = ctx->own_stream;
    {
        int lo = 0, hi = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CU(cudaStreamCreateWithPriority(&ctx->lde_stream, cudaStreamNonBlocking, hi));
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, device) == cudaSuccess && prop.multiProcessorCount > 0) ctx->sm_count = prop.multiProcessorCount;
        else cudaGetLastError();
        const char *e;
        if ((e = getenv("GLP_CO"))) ctx->co_mode = atoi(e);
        if ((e = getenv("GLP_CO_THREADS")) && atoi(e) >= 32 && atoi(e) <= NTT_THREADS) ctx->co_threads = (uint32_t)(atoi(e) / 32) * 32;
        if ((e = getenv("GLP_CO_CTAS")) && atoi(e) >= 1 && atoi(e) <= 8) ctx->co_ctas_per_sm = (uint32_t)atoi(e);
        if ((e = getenv("GLP_CO_COLS")) && atoi(e) >= 8) ctx->co_cols = (uint32_t)(atoi(e) / 8) * 8;
        if ((e = getenv("GLP_CO_FIRST")) && atoi(e) >= 0) ctx->co_first_cols = (uint32_t)(atoi(e) / 8) * 8;
    }
    const char *mq = getenv("GLP_NTT_MAX_LOGQ");
    if (mq && atoi(mq) >= 1 && atoi(mq) <= NTT_MAX_LOGQ) ctx->max_logq = (uint32_t)atoi(mq);
    const char *npt = getenv("GLP_NTT_NO_PASS_TABLES");
    ctx->no_pass_tables = npt && atoi(npt) != 0;
    const char *nfl = getenv("GLP_NTT_NO_FUSED_LDE");
    ctx->no_fused_lde = nfl && atoi(nfl) != 0;
    const char *nn = getenv("GLP_NTT_NO_NARROW");
    ctx->no_narrow = nn && atoi(nn) != 0;
    const char *nsp = getenv("GLP_NTT_NO_SPLIT");
    ctx->no_split = nsp && atoi(nsp) != 0;
    const char *no = getenv("GLP_NTT_OPT");
    if (no) ctx->opt_enabled = atoi(no) != 0;
    const char *nfr = getenv("GLP_NTT_FORCE_RETRY");
    ctx->opt_force_retry = nfr && atoi(nfr) != 0;
    const char *tr = getenv("GLP_TRACE");
    ctx->trace = tr && atoi(tr) != 0;
    const char *msm = getenv("GLP_MERKLE_STREAM_MIN");
    if (msm && atoll(msm) > 0) ctx->merkle_stream_min = (uint64_t)atoll(msm);
    const char *sa = getenv("GLP_STAGE_ALL");
    if (sa) ctx->stage_all = atoi(sa) != 0;
    const char *ds = getenv("GLP_D2H_SPLIT");
    if (ds && atoi(ds) >= 1 && atoi(ds) <= 2) ctx->d2h_split = (uint32_t)atoi(ds);
    const char *sc = getenv("GLP_STREAM_COLS");
    if (sc && atoi(sc) >= 8) ctx->stream_cols = (uint64_t)(atoi(sc) / 8) * 8;
    const char *sg = getenv("GLP_STREAM_GROW");
    if (sg && (atoi(sg) == 0 || atoi(sg) > 100)) ctx->stream_grow = (uint32_t)atoi(sg);
    if (const char *sp = getenv("GLP_STREAM_PLAN"))
        for (const char *q = sp; *q;) {
            char *end = nullptr;
            const long v = strtol(q, &end, 10);
            if (end == q) break;
            if (v >= 8) ctx->stream_plan.push_back((uint64_t)v);
            q = (*end == ',') ? end + 1 : end;
        }
    *out = ctx;
    return GLP_OK;
}

extern "C" void glp_destroy(glp_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (void *p : ctx->tables) cudaFree(p);
    if (ctx->opt_flag_dev) cudaFree(ctx->opt_flag_dev);
    if (ctx->opt_flag_host) cudaFreeHost(ctx->opt_flag_host);
    release(ctx->s1); release(ctx->s2); release(ctx->ext); release(ctx->nodes);
    release(ctx->tmp_in); release(ctx->tmp_out); release(ctx->partial); release(ctx->work); release(ctx->parts_tab);
    cudaStreamDestroy(ctx->own_stream);
    cudaStreamDestroy(ctx->copy_stream[0]);
    cudaStreamDestroy(ctx->copy_stream[1]);
    cudaStreamDestroy(ctx->copy_stream[2]);
    cudaStreamDestroy(ctx->lde_stream);
    delete ctx;
}

extern "C" int glp_set_stream(glp_ctx *ctx, void *s)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    ctx->stream = s ? (cudaStream_t)s : ctx->own_stream;
    return GLP_OK;
}
extern "C" int glp_synchronize(glp_ctx *ctx)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}
extern "C" uint64_t glp_launch_count(const glp_ctx *ctx) { return ctx ? ctx->launches : 0; }
extern "C" uint64_t glp_ntt_retry_count(const glp_ctx *ctx) { return ctx ? ctx->opt_retries : 0; }
extern "C" uint64_t glp_leaf_redo_rows(glp_ctx *ctx)
{
    unsigned long long n = 0;
#if !defined(GLP_EMU)
    if (!ctx || cudaSetDevice(ctx->device) != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess ||
        cudaMemcpyFromSymbol(&n, g_pos_redo_rows, sizeof n) != cudaSuccess) { cudaGetLastError(); return 0; }
#else
    (void)ctx;
    n = __atomic_load_n(&g_pos_redo_rows, __ATOMIC_RELAXED);
#endif
    return n;
}
extern "C" int glp_set_scratch_limit(glp_ctx *ctx, uint64_t bytes)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    ctx->scratch_limit = bytes;
    return GLP_OK;
}

// hand the context's internal device buffers (staging, NTT scratch, private copies of results) back to the driver; they are
// allocated again on demand.  For callers that alternate between the host-pointer and the device-resident entry points
// at sizes where both sets of buffers do not fit next to each other (bench.py's e2e legs at 2^23 x 660).
extern "C" int glp_release_scratch(glp_ctx *ctx)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaStreamSynchronize(ctx->copy_stream[0]));
    CU(cudaStreamSynchronize(ctx->copy_stream[1]));
    CU(cudaStreamSynchronize(ctx->copy_stream[2]));
    release(ctx->s1); release(ctx->s2); release(ctx->ext); release(ctx->nodes);
    release(ctx->tmp_in); release(ctx->tmp_out); release(ctx->work);
    return GLP_OK;
}

extern "C" int glp_device_alloc(glp_ctx *ctx, uint64_t bytes, void **ptr)
{
    if (!ctx || !ptr) return fail(GLP_E_ARG, "glp_device_alloc: NULL argument");
    CU(cudaSetDevice(ctx->device));
    cudaError_t e = cudaMalloc(ptr, bytes ? bytes : 8);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(GLP_E_OOM, std::string("cudaMalloc: ") + cudaGetErrorString(e)); }
    return GLP_OK;
}
extern "C" int glp_device_free(glp_ctx *ctx, void *ptr) { (void)ctx; if (ptr) CU(cudaFree(ptr)); return GLP_OK; }
extern "C" int glp_host_alloc(glp_ctx *ctx, uint64_t bytes, void **ptr)
{
    if (!ctx || !ptr) return fail(GLP_E_ARG, "glp_host_alloc: NULL argument");
    cudaError_t e = cudaHostAlloc(ptr, bytes ? bytes : 8, cudaHostAllocDefault);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(GLP_E_OOM, std::string("cudaHostAlloc: ") + cudaGetErrorString(e)); }
    return GLP_OK;
}
extern "C" int glp_host_free(glp_ctx *ctx, void *ptr) { (void)ctx; if (ptr) CU(cudaFreeHost(ptr)); return GLP_OK; }
// Pin memory the caller already owns (the prover's big polynomial buffer, src/prover/prover.cpp:101-115, is malloc'ed or
// mmap'ed once and reused for every proof): copies to and from it then run at the pinned rate (52-57 GB/s instead of
// 11-21 GB/s pageable, profiles/r01_pcie_probe.txt).  The range stays usable as ordinary memory.
extern "C" int glp_host_register(glp_ctx *ctx, void *ptr, uint64_t bytes)
{
    if (!ctx || !ptr || !bytes) return fail(GLP_E_ARG, "glp_host_register: NULL argument");
    CU(cudaSetDevice(ctx->device));
    CU(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return GLP_OK;
}
extern "C" int glp_host_unregister(glp_ctx *ctx, void *ptr)
{
    if (!ctx || !ptr) return fail(GLP_E_ARG, "glp_host_unregister: NULL argument");
    CU(cudaHostUnregister(ptr));
    return GLP_OK;
}
extern "C" int glp_memcpy(glp_ctx *ctx, void *dst, const void *src, uint64_t bytes)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (!bytes) return GLP_OK;
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}
extern "C" int glp_mem_info(glp_ctx *ctx, uint64_t *free_bytes, uint64_t *total_bytes)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    size_t f = 0, t = 0;
    CU(cudaMemGetInfo(&f, &t));
    if (free_bytes) *free_bytes = f;
    if (total_bytes) *total_bytes = t;
    return GLP_OK;
}

static bool on_device(const void *p)
{
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

static uint32_t ilog2_u64(uint64_t n) { uint32_t r = 0; while (n > 1) { n >>= 1; r++; } return r; }
static bool is_pow2(uint64_t n) { return n && !(n & (n - 1)); }

// ------------------------------------------------------------------------------------------------
// twiddle tables (K1)
// ------------------------------------------------------------------------------------------------
static gl_t host_root(uint32_t k) { return gl_canon(gl_root_of_unity(k)); }

static int table_alloc(glp_ctx *ctx, size_t words, gl_t **out)
{
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, words * sizeof(gl_t));
    if (e != cudaSuccess) { cudaGetLastError(); return fail(GLP_E_OOM, "cudaMalloc(twiddle table) failed"); }
    ctx->tables.push_back(p);
    *out = (gl_t *)p;
    return GLP_OK;
}

// lo[e] = c * g^e (e < 2^min(logN,12)), hi[h] = g^(h << 12) (h < max(1, N >> 12))
static int make_two_level(glp_ctx *ctx, uint32_t logN, gl_t c, gl_t g, TwoLevel *t)
{
    const uint64_t nlo = 1ull << (logN < NTT_LOGT ? logN : NTT_LOGT);
    const uint64_t nhi = logN > NTT_LOGT ? (1ull << (logN - NTT_LOGT)) : 1;
    GLP(table_alloc(ctx, nlo, &t->lo));
    GLP(table_alloc(ctx, nhi, &t->hi));
    LAUNCH_FLAT(ctx, k_ntt_pow_table, (unsigned)((nlo + 127) / 128), 128, t->lo, c, g, (uint64_t)1, nlo);
    LAUNCH_FLAT(ctx, k_ntt_pow_table, (unsigned)((nhi + 127) / 128), 128, t->hi, (gl_t)1, g, (uint64_t)1 << NTT_LOGT, nhi);
    CU(cudaGetLastError());
    return GLP_OK;
}
static int get_tw_tab(glp_ctx *ctx, uint32_t logN, const TwoLevel **out)
{
    auto it = ctx->tw_tabs.find(logN);
    if (it == ctx->tw_tabs.end()) {
        TwoLevel t;
        GLP(make_two_level(ctx, logN, 1, host_root(logN), &t));
        it = ctx->tw_tabs.emplace(logN, t).first;
    }
    *out = &it->second;
    return GLP_OK;
}
static gl_t host_ninv(uint32_t logN) { return gl_canon(gl_inv((gl_t)((1ull << logN) % GL_P))); }
static int get_sc_tab(glp_ctx *ctx, uint32_t logN, const TwoLevel **out)
{
    auto it = ctx->sc_tabs.find(logN);
    if (it == ctx->sc_tabs.end()) {
        TwoLevel t;
        GLP(make_two_level(ctx, logN, host_ninv(logN), GL_SHIFT, &t));
        it = ctx->sc_tabs.emplace(logN, t).first;
    }
    *out = &it->second;
    return GLP_OK;
}
// w_256^i, i < 256: the step-to-step twiddles of every tile size (w_Q^e = w_256^(e * 256/Q))
static int get_zeta(glp_ctx *ctx, const gl_t **out)
{
    if (!ctx->zeta) {
        GLP(table_alloc(ctx, 256, &ctx->zeta));
        LAUNCH_FLAT(ctx, k_ntt_pow_table, 2, 128, ctx->zeta, (gl_t)1, host_root(8), (uint64_t)1, (uint64_t)256);
        CU(cudaGetLastError());
    }
    *out = ctx->zeta;
    return GLP_OK;
}
// Pass table [L/Q][Q] of w_L^(+-p*i): one contiguous row per tile.  NULL (-> two-level tables in the kernel)
// when the pass needs none (L == Q) or the table would exceed the cap.
#define NTT_PASS_TABLE_MAX_LOGL 25
static int get_pass_table(glp_ctx *ctx, uint32_t logL, uint32_t logQ, bool inv, const gl_t **out)
{
    *out = nullptr;
    if (logL <= logQ || logL > NTT_PASS_TABLE_MAX_LOGL || ctx->no_pass_tables) return GLP_OK;
    const uint64_t key = ((uint64_t)logL << 16) | (logQ << 8) | (inv ? 1 : 0);
    auto it = ctx->pass_tabs.find(key);
    if (it == ctx->pass_tabs.end()) {
        gl_t *t = nullptr;
        const uint64_t n = 1ull << logL;
        GLP(table_alloc(ctx, n, &t));
        gl_t g = host_root(logL);
        if (inv) g = gl_canon(gl_inv(g));
        LAUNCH_FLAT(ctx, k_ntt_pass_table, (unsigned)((n + 255) / 256), 256, t, g, logQ, n);
        CU(cudaGetLastError());
        it = ctx->pass_tabs.emplace(key, t).first;
    }
    *out = it->second;
    return GLP_OK;
}

extern "C" uint64_t glp_root(uint32_t domain_pow, uint64_t idx)
{
    if (domain_pow > 32) return 0;
    return gl_canon(gl_pow(host_root(domain_pow), idx));
}

// ------------------------------------------------------------------------------------------------
// NTT passes (K2-K5)
// ------------------------------------------------------------------------------------------------
static bool plan_passes(const glp_ctx *ctx, uint32_t logN, NttPlanLogs *pl)
{
    const uint32_t mq = ctx->max_logq;
    memset(pl, 0, sizeof *pl);
    if (logN == 0) return true;
    const uint32_t np = (logN + mq - 1) / mq;
    if (np > NTT_MAX_PASSES) return false;
    pl->npass = np;
    for (uint32_t i = 0; i < np; i++) pl->lg[i] = logN / np + (i < logN % np ? 1 : 0);
    return true;
}
// shift^j/N in the storage order the DIF plan `pl` leaves coefficient j in
static int get_scale_table(glp_ctx *ctx, uint32_t logN, const NttPlanLogs &pl, const gl_t **out, bool masked = false)
{
    *out = nullptr;
    if (logN > NTT_PASS_TABLE_MAX_LOGL || ctx->no_pass_tables) return GLP_OK;
    uint64_t key = logN | (masked ? 0x80u : 0u);                  // masked: shift^j / (N * NTT_MASK), for the optimistic LDE
    for (uint32_t i = 0; i < NTT_MAX_PASSES; i++) key = (key << 8) | (i < pl.npass ? pl.lg[i] : 0);
    auto it = ctx->scale_tabs.find(key);
    if (it == ctx->scale_tabs.end()) {
        gl_t *t = nullptr;
        const uint64_t n = 1ull << logN;
        GLP(table_alloc(ctx, n, &t));
        const gl_t c = masked ? gl_canon(gl_mul(host_ninv(logN), gl_inv((gl_t)NTT_MASK))) : host_ninv(logN);
        LAUNCH_FLAT(ctx, k_ntt_scale_table, (unsigned)((n + 255) / 256), 256, t, c, (gl_t)GL_SHIFT, pl, n);
        CU(cudaGetLastError());
        it = ctx->scale_tabs.emplace(key, t).first;
    }
    *out = it->second;
    return GLP_OK;
}

#define NTT_NARROW_MAX_COLS 16
template <int A, int B>
static int launch_pass_ab(glp_ctx *ctx, const NttPass &a_in, bool inv, uint32_t lsh)
{
    // A matrix of 32 k + r columns leaves 32 - r lanes of its last column group idle (a rank's slice of 660 columns on 4 / 8
    // devices is 165 / 80-88 columns wide: 5 / 16 / 24 columns in the last group).  The remainder runs as its own launch on the
    // NARROW mapping instead -- 2^log_cw >= r lanes per tile, 32 >> log_cw row tiles per warp -- when that takes fewer warps:
    // r <= 16 in one launch; 16 < r <= 24 as 16 columns plus r - 16 <= 8 columns (1/2 + 1/4 of a group).  The launches touch
    // disjoint columns; passes are in place per column, so their order does not matter.  GLP_NTT_NO_SPLIT=1 turns it off.
    if (!ctx->no_narrow && !ctx->no_split && !ctx->co_active && a_in.ncols > NTT_LANES) {
        const uint32_t r = a_in.ncols % NTT_LANES;
        uint32_t tail = 0;                                        // columns handed to a launch of their own
        if (r && r <= NTT_NARROW_MAX_COLS) tail = r;
        else if (r > 16 && r <= 24) tail = r - 16;
        if (tail) {
            NttPass head = a_in, rest = a_in;
            head.ncols = a_in.ncols - tail;
            rest.ncols = tail;
            rest.in += head.ncols;
            rest.out += head.ncols;
            rest.part_col0 += head.ncols;
            GLP((launch_pass_ab<A, B>(ctx, head, inv, lsh)));
            return launch_pass_ab<A, B>(ctx, rest, inv, lsh);
        }
    }
    NttPass a = a_in;
    const bool narrow = a.ncols <= NTT_NARROW_MAX_COLS && !ctx->no_narrow;
    const uint64_t ntiles = 1ull << (a.logN - (A + B));
    uint64_t grid;
    if (narrow) {
        a.log_cw = 0;
        while ((1u << a.log_cw) < a.ncols) a.log_cw++;
        const uint64_t tiles_per_warp = 32u >> a.log_cw;
        const uint64_t groups = (ntiles + tiles_per_warp - 1) / tiles_per_warp;          // warps' worth of tiles
        grid = B == 0 ? (groups + NTT_WARPS - 1) / NTT_WARPS : groups;
    } else {
        const uint64_t ncg = (a.ncols + NTT_LANES - 1) / NTT_LANES;
        grid = (B == 0 ? (ntiles + NTT_WARPS - 1) / NTT_WARPS : ntiles) * ncg;
    }
    if (grid > 0x7FFFFFFFull) return fail(GLP_E_ARG, "NTT grid too large");
    const size_t smem = B == 0 ? 0 : ((size_t)1 << (A + B)) * NTT_LANES * sizeof(gl_t);
#define NTT_GO2(INV, LSH, NARROW)                                                    \
    do {                                                                             \
        auto kfn = k_ntt_pass<A, B, INV, LSH, NARROW>;                               \
        if (smem > 48 * 1024) GLP_SET_SMEM(kfn, smem);                               \
        LAUNCH(ctx, kfn, (unsigned)grid, NTT_THREADS, smem, a);                      \
    } while (0)
#define NTT_GO_CO(INV, LSH)                                                          \
    do {                                                                             \
        if constexpr (B > 0 && A + B >= 7) {                                         \
            auto kfn = k_ntt_pass_co<A, B, INV, LSH>;                                \
            if (smem > 48 * 1024) GLP_SET_SMEM(kfn, smem);                           \
            const uint64_t cog = (uint64_t)ctx->sm_count * ctx->co_ctas_per_sm;      \
            LAUNCH(ctx, kfn, (unsigned)(grid < cog ? grid : cog), ctx->co_threads, smem, a, (uint32_t)grid); \
        }                                                                            \
    } while (0)
#define NTT_GO_OPT(INV, LSH)                                                         \
    do {                                                                             \
        if constexpr (B > 0 && A + B >= 6) {                                         \
            auto kfn = k_ntt_pass_opt<A, B, INV, LSH>;                               \
            if (smem > 48 * 1024) GLP_SET_SMEM(kfn, smem);                           \
            a.defer_flag = ctx->opt_flag_dev;                                        \
            LAUNCH(ctx, kfn, (unsigned)grid, NTT_THREADS, smem, a);                  \
        }                                                                            \
    } while (0)
#define NTT_GO(INV, LSH)                                                             \
    do {                                                                             \
        if (narrow) NTT_GO2(INV, LSH, true);                                         \
        else if (B > 0 && A + B >= 7 && ctx->co_active) NTT_GO_CO(INV, LSH);         \
        else if (B > 0 && A + B >= 6 && a.opt) NTT_GO_OPT(INV, LSH);       \
        else NTT_GO2(INV, LSH, false);                                               \
    } while (0)
    if (inv) NTT_GO(true, 0);
    else if (A == 4 && lsh == 1) NTT_GO(false, (A == 4 ? 1 : 0));
    else if (A == 4 && lsh == 2) NTT_GO(false, (A == 4 ? 2 : 0));
    else if (A == 4 && lsh == 3) NTT_GO(false, (A == 4 ? 3 : 0));
    else NTT_GO(false, 0);
#undef NTT_GO
#undef NTT_GO_CO
#undef NTT_GO_OPT
#undef NTT_GO2
    return GLP_OK;
}
static int launch_pass(glp_ctx *ctx, const NttPass &a, uint32_t lg, bool inv, uint32_t lsh)
{
    switch (lg) {
    case 8: return launch_pass_ab<4, 4>(ctx, a, inv, lsh);
    case 7: return launch_pass_ab<4, 3>(ctx, a, inv, lsh);
    case 6: return launch_pass_ab<3, 3>(ctx, a, inv, 0);
    case 5: return launch_pass_ab<3, 2>(ctx, a, inv, 0);
    case 4: return launch_pass_ab<4, 0>(ctx, a, inv, lsh);
    case 3: return launch_pass_ab<3, 0>(ctx, a, inv, 0);
    case 2: return launch_pass_ab<2, 0>(ctx, a, inv, 0);
    case 1: return launch_pass_ab<1, 0>(ctx, a, inv, 0);
    }
    return fail(GLP_E_ARG, "NTT: bad tile size");
}

static int copy_canon(glp_ctx *ctx, gl_t *out, uint64_t out_pitch, const gl_t *in, uint64_t in_pitch, uint64_t nrows, uint64_t ncols)
{
    const uint64_t nb = (nrows * ncols + 255) / 256;
    LAUNCH_FLAT(ctx, k_ntt_copy_canon, (unsigned)(nb > 148 * 32 ? 148 * 32 : nb), 256, out, out_pitch, in, in_pitch, nrows, ncols);
    CU(cudaGetLastError());
    return GLP_OK;
}

// fields common to every pass of one transform
static int pass_common(glp_ctx *ctx, NttPass *a, uint32_t logN, uint32_t wb, bool need_two_level)
{
    memset(a, 0, sizeof *a);
    a->ncols = wb;
    a->logN = logN;
    GLP(get_zeta(ctx, &a->zeta));
    if (need_two_level) {
        const TwoLevel *tw = nullptr;
        GLP(get_tw_tab(ctx, logN, &tw));
        a->tw_lo = tw->lo;
        a->tw_hi = tw->hi;
    }
    return GLP_OK;
}

// Natural order in, natural order out, `wb` columns: in (pitch) -> out (pitch); scratch holds 2^logN x wb words
// (needed when the plan has more than one pass).  Rows >= live of `in` are zero and never read.
static int run_transform(glp_ctx *ctx, uint32_t logN, const gl_t *in, uint64_t in_pitch, gl_t *out, uint64_t out_pitch, uint32_t wb,
                         uint64_t live, bool inverse, bool extend, gl_t *scratch)
{
    if (logN == 0) return copy_canon(ctx, out, out_pitch, in, in_pitch, 1, wb);     // 1/N = shift^0 = 1
    NttPlanLogs pl;
    if (!plan_passes(ctx, logN, &pl)) return fail(GLP_E_ARG, "NTT: size exceeds four tile passes");
    uint32_t logL = logN;
    for (uint32_t i = 0; i < pl.npass; i++) {
        const bool first = i == 0, last = i + 1 == pl.npass;
        NttPass a;
        const gl_t *tab = nullptr;
        if (!last) GLP(get_pass_table(ctx, logL, pl.lg[i], inverse, &tab));
        GLP(pass_common(ctx, &a, logN, wb, !last && !tab));
        a.in = first ? in : scratch;
        a.in_pitch = first ? in_pitch : wb;
        a.out = last ? out : scratch;
        a.out_pitch = last ? out_pitch : wb;
        a.logL = a.log_in_L = logL;
        a.live = first ? live : (1ull << logL);
        a.tw_mode = last ? NTT_TW_NONE : NTT_TW_POST;
        a.tw = tab;
        a.scatter = last;
        a.nprev = i;
        for (uint32_t k = 0; k < i; k++) a.prev_log[k] = pl.lg[k];
        a.canon_in = first;
        a.canon_out = last;
        if (last && inverse) {
            if (extend) {
                const TwoLevel *sc = nullptr;
                GLP(get_sc_tab(ctx, logN, &sc));
                a.scale_mode = NTT_SC_TWOLEVEL;
                a.sc_lo = sc->lo;
                a.sc_hi = sc->hi;
            } else {
                a.scale_mode = NTT_SC_CONST;
                a.scale_const = host_ninv(logN);
            }
        }
        // compile-time pruning when exactly the upper 1 - 2^-lsh of the first pass's tile rows are zero
        uint32_t lsh = 0;                                         // rows q*S + p < N >> e  <=>  high digit of q < 16 >> e
        if (first && !inverse)
            for (uint32_t e = 1; e <= 3 && e <= logN; e++)
                if (live == (1ull << (logN - e))) lsh = e;
        GLP(launch_pass(ctx, a, pl.lg[i], inverse, lsh));
        logL -= pl.lg[i];
    }
    CU(cudaGetLastError());
    return GLP_OK;
}

// Low-degree extension of `wb` columns without any reordering pass: INTT decimation-in-frequency (natural in,
// digit-reversed out, in place in scratch S = 2^logN x wb), coset scaling on its last pass, then the size-2^(logN+e)
// forward NTT decimation-in-time with the same radices, reading S and finishing in place in `out`.
// Returns 1 when the shape has no such plan (caller falls back to two natural-order transforms).
// where the rows of an extension go when it is row-sharded over several slabs (multi-GPU: one slab per rank)
struct ScatterDst {
    const unsigned long long *parts;         // device array of nparts slab bases
    uint32_t nparts, log_part_rows;
    uint64_t pitch, col0;
};

static int run_lde(glp_ctx *ctx, uint32_t logN, uint32_t e, const gl_t *in, uint64_t in_pitch, gl_t *out, uint64_t out_pitch, uint32_t wb,
                   gl_t *S, const ScatterDst *sc = nullptr)
{
    const uint32_t logM = logN + e;
    NttPlanLogs q, r;
    if (!plan_passes(ctx, logM, &q) || q.npass == 0) return 1;
    if (q.lg[q.npass - 1] < e) return 1;
    r = q;
    r.lg[r.npass - 1] -= e;
    if (r.lg[r.npass - 1] == 0) r.npass--;
    bool opt = false;
    // ---- INTT, DIF, in place in S ----
    {
        ProfScope ps(ctx, 0);
        if (r.npass == 0) GLP(copy_canon(ctx, S, wb, in, in_pitch, 1, wb));
        const gl_t *sct = nullptr;
        if (r.npass) GLP(get_scale_table(ctx, logN, r, &sct));
        if (ctx->opt_active && sct) {                            // optimistic: masked inputs, mask undone by the scale table
            GLP(get_scale_table(ctx, logN, r, &sct, true));
            opt = true;
        }
        uint32_t logL = logN;
        for (uint32_t i = 0; i < r.npass; i++) {
            const bool first = i == 0, last = i + 1 == r.npass;
            NttPass a;
            const gl_t *tab = nullptr;
            if (!last) GLP(get_pass_table(ctx, logL, r.lg[i], true, &tab));
            GLP(pass_common(ctx, &a, logN, wb, !last && !tab));
            a.in = first ? in : S;
            a.in_pitch = first ? in_pitch : wb;
            a.out = S;
            a.out_pitch = wb;
            a.logL = a.log_in_L = logL;
            a.live = 1ull << logL;
            a.tw_mode = last ? NTT_TW_NONE : NTT_TW_POST;
            a.tw = tab;
            a.nprev = i;
            for (uint32_t k = 0; k < i; k++) a.prev_log[k] = r.lg[k];
            a.canon_in = first;
            a.opt = opt;
            if (first && opt) a.mask = (gl_t)NTT_MASK;
            if (last) {
                if (sct) { a.scale_mode = NTT_SC_TABLE; a.sc = sct; }
                else {
                    const TwoLevel *sc = nullptr;
                    GLP(get_sc_tab(ctx, logN, &sc));
                    a.scale_mode = NTT_SC_TWOLEVEL;
                    a.sc_lo = sc->lo;
                    a.sc_hi = sc->hi;
                }
            }
            GLP(launch_pass(ctx, a, r.lg[i], true, 0));
            logL -= r.lg[i];
        }
    }
    // ---- forward NTT of size 2^logM, DIT: last digit (contiguous rows) first ----
    {
        ProfScope ps(ctx, 1);
        uint32_t logL = 0;
        for (int i = (int)q.npass - 1; i >= 0; i--) {
            logL += q.lg[i];
            const bool first = i == (int)q.npass - 1, last = i == 0;
            NttPass a;
            const gl_t *tab = nullptr;
            if (!first) GLP(get_pass_table(ctx, logL, q.lg[i], false, &tab));
            GLP(pass_common(ctx, &a, logM, wb, !first && !tab));
            a.in = first ? S : out;
            a.in_pitch = first ? wb : out_pitch;
            a.out = out;
            a.out_pitch = out_pitch;
            a.logL = logL;
            a.log_in_L = first ? logL - e : logL;
            a.live = first ? (1ull << (logL - e)) : (1ull << logL);
            a.tw_mode = first ? NTT_TW_NONE : NTT_TW_PRE;
            a.tw = tab;
            a.canon_out = last;
            a.opt = opt;
            if (last && sc) {                                     // the exchange rides on the last pass's stores
                a.parts = sc->parts;
                a.nparts = sc->nparts;
                a.log_part_rows = sc->log_part_rows;
                a.part_pitch = sc->pitch;
                a.part_col0 = sc->col0;
            }
            uint32_t lsh = 0;
            if (first && e >= 1 && e <= 3) lsh = e;              // used by the kernels whose high digit of q is 4 bits wide
            GLP(launch_pass(ctx, a, q.lg[i], false, lsh));
        }
    }
    CU(cudaGetLastError());
    return GLP_OK;
}

static uint64_t scratch_budget(glp_ctx *ctx)
{
    if (ctx->scratch_limit) return ctx->scratch_limit;
    size_t f = 0, t = 0;
    if (cudaMemGetInfo(&f, &t) != cudaSuccess) { cudaGetLastError(); f = (size_t)1 << 30; }
    uint64_t b = (uint64_t)f / 4 + ctx->s1.bytes + ctx->s2.bytes;
    const uint64_t cap = 16ull << 30;
    return b > cap ? cap : b;
}
// columns per block so that `rows_total` scratch rows fit the budget
static uint32_t pick_block_cols(glp_ctx *ctx, uint64_t rows_total, uint64_t ncols)
{
    uint64_t wb = scratch_budget(ctx) / (rows_total * sizeof(gl_t));
    if (wb >= ncols) return (uint32_t)ncols;
    if (wb >= 32) wb = 32 * (wb / 32);
    if (wb < 1) wb = 1;
    return (uint32_t)wb;
}

// device pointers only
static int dev_ntt(glp_ctx *ctx, gl_t *dst, uint64_t dst_pitch, const gl_t *src, uint64_t src_pitch, uint64_t size, uint64_t ncols,
                   bool inverse, bool extend, uint64_t live = 0)
{
    if (!live || live > size) live = size;
    const uint32_t logN = ilog2_u64(size);
    NttPlanLogs pl;
    if (!plan_passes(ctx, logN, &pl)) return fail(GLP_E_ARG, "NTT: size exceeds four tile passes");
    uint32_t wb = (uint32_t)ncols;
    if (pl.npass > 1) {
        wb = pick_block_cols(ctx, size, ncols);
        GLP(ensure(ctx, ctx->s1, (size_t)size * wb * sizeof(gl_t)));
    }
    for (uint64_t c0 = 0; c0 < ncols; c0 += wb) {
        const uint32_t w = (uint32_t)((ncols - c0 < wb) ? ncols - c0 : wb);
        GLP(run_transform(ctx, logN, src + c0, src_pitch, dst + c0, dst_pitch, w, live, inverse, extend, (gl_t *)ctx->s1.p));
    }
    return GLP_OK;
}

// one column block of the extension: in (n rows) -> out (n_ext rows); S holds n x w words, S2 n_ext x w (fallback only)
static int lde_block(glp_ctx *ctx, uint32_t logN, uint32_t logE, const gl_t *in, uint64_t in_pitch, gl_t *out, uint64_t out_pitch, uint32_t w,
                     const ScatterDst *sc = nullptr)
{
    const uint64_t n = 1ull << logN, n_ext = 1ull << logE;
    GLP(ensure(ctx, ctx->s2, (size_t)n * w * sizeof(gl_t)));
    int rc = ctx->no_fused_lde ? 1 : run_lde(ctx, logN, logE - logN, in, in_pitch, out, out_pitch, w, (gl_t *)ctx->s2.p, sc);
    if (rc != 1) return rc;
    if (sc) {                                                     // no fused plan for this shape: extend into `out`, then scatter its rows
        GLP(lde_block(ctx, logN, logE, in, in_pitch, out, out_pitch, w, nullptr));
        const uint64_t nb = (n_ext * w + 255) / 256;
        LAUNCH_FLAT(ctx, k_ntt_scatter_rows, (unsigned)(nb > 148 * 32 ? 148 * 32 : nb), 256, sc->parts, sc->log_part_rows, sc->pitch, sc->col0,
                    (const gl_t *)out, out_pitch, n_ext, (uint64_t)w);
        CU(cudaGetLastError());
        return GLP_OK;
    }
    // fallback: two natural-order transforms (INTT with coset scaling, then the zero-padded forward NTT)
    NttPlanLogs pe;
    if (!plan_passes(ctx, logE, &pe)) return fail(GLP_E_ARG, "NTT: size exceeds four tile passes");
    GLP(ensure(ctx, ctx->s1, (size_t)n_ext * w * sizeof(gl_t)));
    {
        ProfScope ps(ctx, 0);
        GLP(run_transform(ctx, logN, in, in_pitch, (gl_t *)ctx->s2.p, w, w, n, true, true, (gl_t *)ctx->s1.p));
    }
    {
        ProfScope ps(ctx, 1);
        GLP(run_transform(ctx, logE, (gl_t *)ctx->s2.p, w, out, out_pitch, w, n, false, false, (gl_t *)ctx->s1.p));
    }
    return GLP_OK;
}

static int dev_extend_pol(glp_ctx *ctx, gl_t *out, uint64_t out_pitch, const gl_t *in, uint64_t in_pitch, uint64_t n_ext, uint64_t n,
                          uint64_t ncols)
{
    const uint32_t logN = ilog2_u64(n), logE = ilog2_u64(n_ext);
    const uint32_t wb = pick_block_cols(ctx, ctx->no_fused_lde ? n_ext + n : n, ncols);
    for (uint64_t c0 = 0; c0 < ncols; c0 += wb) {
        const uint32_t w = (uint32_t)((ncols - c0 < wb) ? ncols - c0 : wb);
        GLP(lde_block(ctx, logN, logE, in + c0, in_pitch, out + c0, out_pitch, w));
    }
    return GLP_OK;
}

// Optimistic execution of a device-resident LDE: `issue` enqueues the whole transform on ctx->stream.  It is enqueued first
// with the passes whose butterflies leave out the second wrap of every add / subtract (k_ntt_pass_opt; the wrap fires with
// probability 2^-64 per operation on honest data), and -- only if one of them raised the flag, which the host reads after a
// stream synchronise -- once more with the exact passes.  The input must survive the first attempt (`rerunnable`: no
// overlap between input and output), otherwise the exact passes run directly.  Returns with the stream idle.
template <class F>
static int with_optimistic_ntt(glp_ctx *ctx, bool rerunnable, F &&issue)
{
    if (!ctx->opt_enabled || !rerunnable || ctx->no_fused_lde) {
        GLP(issue());
        CU(cudaStreamSynchronize(ctx->stream));
        return GLP_OK;
    }
    if (!ctx->opt_flag_dev) {
        CU(cudaMalloc((void **)&ctx->opt_flag_dev, sizeof(uint32_t)));
        CU(cudaHostAlloc((void **)&ctx->opt_flag_host, sizeof(uint32_t), cudaHostAllocDefault));
    }
    CU(cudaMemsetAsync(ctx->opt_flag_dev, 0, sizeof(uint32_t), ctx->stream));
    ctx->opt_active = true;
    const int rc = issue();
    ctx->opt_active = false;
    if (rc != GLP_OK) return rc;
    CU(cudaMemcpyAsync(ctx->opt_flag_host, ctx->opt_flag_dev, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (*ctx->opt_flag_host || ctx->opt_force_retry) {
        ctx->opt_retries++;
        GLP(issue());
        CU(cudaStreamSynchronize(ctx->stream));
    }
    return GLP_OK;
}
static int dev_extend_pol_checked(glp_ctx *ctx, gl_t *out, uint64_t out_pitch, const gl_t *in, uint64_t in_pitch, uint64_t n_ext, uint64_t n,
                                  uint64_t ncols)
{
    const gl_t *in_end = in + (n - 1) * in_pitch + ncols, *out_end = out + (n_ext - 1) * out_pitch + ncols;
    return with_optimistic_ntt(ctx, in_end <= out || out_end <= in,
                               [&]() -> int { return dev_extend_pol(ctx, out, out_pitch, in, in_pitch, n_ext, n, ncols); });
}

// ------------------------------------------------------------------------------------------------
// host <-> device staging helpers
// ------------------------------------------------------------------------------------------------
// Bring a host matrix to a private device buffer (or pass a device pointer through).
static int stage_in(glp_ctx *ctx, DevBuf &buf, const uint64_t *p, size_t words, const gl_t **dev)
{
    if (on_device(p)) { *dev = (const gl_t *)p; return GLP_OK; }
    GLP(ensure(ctx, buf, words * sizeof(gl_t)));
    CU(cudaMemcpyAsync(buf.p, p, words * sizeof(gl_t), cudaMemcpyHostToDevice, ctx->stream));
    *dev = (const gl_t *)buf.p;
    return GLP_OK;
}

// ------------------------------------------------------------------------------------------------
// NTT entry points
// ------------------------------------------------------------------------------------------------
extern "C" int glp_ntt(glp_ctx *ctx, uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, int inverse, int extend)
{
    return glp_ntt_live(ctx, dst, src, size, ncols, inverse, extend, size);
}
extern "C" int glp_ntt_live(glp_ctx *ctx, uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, int inverse, int extend,
                            uint64_t live_rows)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (size == 0 || ncols == 0) return GLP_OK;                 // ntt_goldilocks.cpp:168-171
    if (!src) return fail(GLP_E_ARG, "glp_ntt: src is NULL");
    if (!is_pow2(size)) return fail(GLP_E_ARG, "glp_ntt: size is not a power of two");
    if (size > (1ull << 32)) return fail(GLP_E_DOMAIN, "Domain size too big for the curve");
    if (ilog2_u64(size) > NTT_MAX_PASSES * ctx->max_logq) return fail(GLP_E_ARG, "glp_ntt: size exceeds four tile passes");
    if (!dst) dst = (uint64_t *)src;                            // ntt_goldilocks.cpp:15-22
    CU(cudaSetDevice(ctx->device));
    const size_t words = (size_t)size * ncols;
    const bool dst_dev = on_device(dst);
    const gl_t *dsrc = nullptr;
    GLP(stage_in(ctx, ctx->tmp_in, src, words, &dsrc));
    gl_t *ddst = (gl_t *)dst;
    if (!dst_dev) {
        // run in place on the staged copy when the source was staged, else into a private buffer
        if ((const void *)dsrc == ctx->tmp_in.p) ddst = (gl_t *)ctx->tmp_in.p;
        else { GLP(ensure(ctx, ctx->tmp_out, words * sizeof(gl_t))); ddst = (gl_t *)ctx->tmp_out.p; }
    }
    if (live_rows == 0) {                                        // everything is read as zero: the transform of zero
        CU(cudaMemsetAsync(ddst, 0, words * sizeof(gl_t), ctx->stream));
    } else {
        GLP(dev_ntt(ctx, ddst, ncols, dsrc, ncols, size, ncols, inverse != 0, inverse != 0 && extend != 0, live_rows));
    }
    if (!dst_dev) CU(cudaMemcpyAsync(dst, ddst, words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

extern "C" int glp_reverse_permutation(glp_ctx *ctx, uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t offset_cols,
                                       uint64_t ncols, uint64_t ncols_all, uint64_t extension)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (size == 0 || ncols == 0) return GLP_OK;
    if (!dst || !src) return fail(GLP_E_ARG, "glp_reverse_permutation: NULL pointer");
    if (!is_pow2(size) || size > (1ull << 32)) return fail(GLP_E_ARG, "glp_reverse_permutation: size must be a power of two <= 2^32");
    if (offset_cols + ncols > ncols_all) return fail(GLP_E_ARG, "glp_reverse_permutation: column window out of range");
    if (dst == src && (extension > 1 || ncols != ncols_all || offset_cols != 0))
        return fail(GLP_E_ARG, "glp_reverse_permutation: in place needs a single block and extension <= 1");   // :280-304
    CU(cudaSetDevice(ctx->device));
    const uint64_t live = extension > 1 ? size / extension : size;
    // only rows < live of src are ever read
    const size_t src_words = (size_t)(live - 1) * ncols_all + offset_cols + ncols, dst_words = (size_t)size * ncols;
    const gl_t *dsrc = nullptr;
    GLP(stage_in(ctx, ctx->tmp_in, src, src_words, &dsrc));
    const bool dst_dev = on_device(dst);
    gl_t *ddst = (gl_t *)dst;
    if (!dst_dev || (const void *)ddst == (const void *)dsrc) { GLP(ensure(ctx, ctx->tmp_out, dst_words * sizeof(gl_t))); ddst = (gl_t *)ctx->tmp_out.p; }
    const uint64_t nb = (dst_words + 255) / 256;
    LAUNCH_FLAT(ctx, k_reverse_permutation, (unsigned)(nb > 148 * 32 ? 148 * 32 : nb), 256, ddst, dsrc, size, ilog2_u64(size), offset_cols,
                ncols, ncols_all, live);
    CU(cudaGetLastError());
    if ((void *)ddst != (void *)dst) CU(cudaMemcpyAsync(dst, ddst, dst_words * sizeof(gl_t), cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

static int check_extend_args(const void *out, const void *in, uint64_t n_ext, uint64_t n)
{
    if (!in || !out) return fail(GLP_E_ARG, "extend_pol: NULL pointer");
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GLP_E_ARG, "extend_pol: sizes must be powers of two with n_ext >= n");
    if (n_ext > (1ull << 32)) return fail(GLP_E_DOMAIN, "Domain size too big for the curve");
    return GLP_OK;
}

extern "C" int glp_extend_pol_pitched(glp_ctx *ctx, uint64_t *out, uint64_t out_pitch, const uint64_t *in, uint64_t in_pitch,
                                      uint64_t n_ext, uint64_t n, uint64_t ncols)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (n == 0 || ncols == 0) return GLP_OK;
    GLP(check_extend_args(out, in, n_ext, n));
    if (in_pitch < ncols || out_pitch < ncols) return fail(GLP_E_ARG, "extend_pol: pitch smaller than ncols");
    CU(cudaSetDevice(ctx->device));
    if (!on_device(in) || !on_device(out)) return fail(GLP_E_ARG, "glp_extend_pol_pitched: device pointers required");
    GLP(dev_extend_pol(ctx, (gl_t *)out, out_pitch, (const gl_t *)in, in_pitch, n_ext, n, ncols));
    return GLP_OK;
}

extern "C" int glp_extend_pol_scatter(glp_ctx *ctx, uint64_t *const *parts, uint32_t nparts, uint64_t part_pitch, uint64_t part_col0,
                                      const uint64_t *in, uint64_t in_pitch, uint64_t n_ext, uint64_t n, uint64_t ncols)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (n == 0 || ncols == 0) return GLP_OK;
    if (!parts || !in) return fail(GLP_E_ARG, "glp_extend_pol_scatter: NULL pointer");
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GLP_E_ARG, "extend_pol: sizes must be powers of two with n_ext >= n");
    if (n_ext > (1ull << 32)) return fail(GLP_E_DOMAIN, "Domain size too big for the curve");
    if (!is_pow2(nparts) || nparts > n_ext || nparts > 64) return fail(GLP_E_ARG, "glp_extend_pol_scatter: nparts must be a power of two <= min(n_ext, 64)");
    if (in_pitch < ncols || part_pitch < part_col0 + ncols) return fail(GLP_E_ARG, "glp_extend_pol_scatter: pitch too small");
    CU(cudaSetDevice(ctx->device));
    for (uint32_t i = 0; i < nparts; i++)
        if (!parts[i]) return fail(GLP_E_ARG, "glp_extend_pol_scatter: NULL slab");
    GLP(ensure(ctx, ctx->parts_tab, 64 * sizeof(unsigned long long)));
    CU(cudaMemcpyAsync(ctx->parts_tab.p, parts, nparts * sizeof(unsigned long long), cudaMemcpyHostToDevice, ctx->stream));
    const uint32_t logN = ilog2_u64(n), logE = ilog2_u64(n_ext);
    ScatterDst sc = {(const unsigned long long *)ctx->parts_tab.p, nparts, logE - ilog2_u64(nparts), part_pitch, part_col0};
    if (!on_device(in)) {
        // host input: upload column block k+1 (strided 2-D copy) while block k is extended and scattered
        uint64_t w = ctx->stream_cols;
        if (w > ncols) w = ncols;
        uint64_t nblocks = (ncols + w - 1) / w;
        if (nblocks == 1 && ncols >= 64) nblocks = 2;              // something to overlap the first upload with
        w = ((ncols + nblocks - 1) / nblocks + 7) / 8 * 8;         // equal blocks (a 165-column slice is 88 + 77, not 160 + 5)
        nblocks = (ncols + w - 1) / w;
        GLP(ensure(ctx, ctx->tmp_in, 2 * (size_t)n * w * sizeof(gl_t)));
        GLP(ensure(ctx, ctx->work, (size_t)n_ext * w * sizeof(gl_t)));
        gl_t *stage[2] = {(gl_t *)ctx->tmp_in.p, (gl_t *)ctx->tmp_in.p + (size_t)n * w};
        std::vector<cudaEvent_t> up(nblocks), used(nblocks);
        PipelineEvents pe(ctx);
        for (uint64_t k = 0; k < nblocks; k++) {
            GLP(pe.make(&up[k], cudaEventDisableTiming));
            GLP(pe.make(&used[k], cudaEventDisableTiming));
        }
        cudaEvent_t start;
        GLP(pe.make(&start, cudaEventDisableTiming));
        CU(cudaEventRecord(start, ctx->stream));
        CU(cudaStreamWaitEvent(ctx->copy_stream[0], start, 0));
        auto width = [&](uint64_t k) { return (uint32_t)((ncols - k * w < w) ? ncols - k * w : w); };
        auto upload = [&](uint64_t k) -> int {
            if (k >= 2) CU(cudaStreamWaitEvent(ctx->copy_stream[0], used[k - 2], 0));
            CU(cudaMemcpy2DAsync(stage[k & 1], (size_t)width(k) * 8, in + k * w, (size_t)in_pitch * 8, (size_t)width(k) * 8, n,
                                 cudaMemcpyHostToDevice, ctx->copy_stream[0]));
            CU(cudaEventRecord(up[k], ctx->copy_stream[0]));
            return GLP_OK;
        };
        GLP(upload(0));
        if (nblocks > 1) GLP(upload(1));
        for (uint64_t k = 0; k < nblocks; k++) {
            CU(cudaStreamWaitEvent(ctx->stream, up[k], 0));
            sc.col0 = part_col0 + k * w;
            GLP(lde_block(ctx, logN, logE, stage[k & 1], width(k), (gl_t *)ctx->work.p, width(k), width(k), &sc));
            CU(cudaEventRecord(used[k], ctx->stream));
            if (k + 2 < nblocks) GLP(upload(k + 2));
        }
        CU(cudaStreamSynchronize(ctx->stream));
        CU(cudaStreamSynchronize(ctx->copy_stream[0]));
        return GLP_OK;
    }
    // device input (which must not overlap any slab): optimistic passes, repeated exactly if they flag; returns synchronised
    const uint32_t wb = pick_block_cols(ctx, n + n_ext, ncols);
    GLP(ensure(ctx, ctx->work, (size_t)n_ext * wb * sizeof(gl_t)));
    return with_optimistic_ntt(ctx, true, [&]() -> int {
        for (uint64_t c0 = 0; c0 < ncols; c0 += wb) {
            const uint32_t w = (uint32_t)((ncols - c0 < wb) ? ncols - c0 : wb);
            sc.col0 = part_col0 + c0;
            GLP(lde_block(ctx, logN, logE, (const gl_t *)in + c0, in_pitch, (gl_t *)ctx->work.p, w, w, &sc));
        }
        return GLP_OK;
    });
}

extern "C" int glp_ipc_export(glp_ctx *ctx, void *devptr, unsigned char handle[64])
{
    if (!ctx || !devptr || !handle) return fail(GLP_E_ARG, "glp_ipc_export: NULL argument");
    CU(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    static_assert(sizeof h == 64, "cudaIpcMemHandle_t is 64 bytes");
    CU(cudaIpcGetMemHandle(&h, devptr));
    memcpy(handle, &h, 64);
    return GLP_OK;
}
extern "C" int glp_ipc_import(glp_ctx *ctx, const unsigned char handle[64], void **devptr)
{
    if (!ctx || !devptr || !handle) return fail(GLP_E_ARG, "glp_ipc_import: NULL argument");
    CU(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    CU(cudaIpcOpenMemHandle(devptr, h, cudaIpcMemLazyEnablePeerAccess));
    return GLP_OK;
}
extern "C" int glp_ipc_close(glp_ctx *ctx, void *devptr)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (devptr) CU(cudaIpcCloseMemHandle(devptr));
    return GLP_OK;
}

extern "C" int glp_extend_pol(glp_ctx *ctx, uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (n == 0 || ncols == 0) return GLP_OK;
    GLP(check_extend_args(out, in, n_ext, n));
    CU(cudaSetDevice(ctx->device));
    const bool in_dev = on_device(in), out_dev = on_device(out);
    if (in_dev && out_dev) {
        return dev_extend_pol_checked(ctx, (gl_t *)out, ncols, (const gl_t *)in, ncols, n_ext, n, ncols);
    }
    return glp_commit(ctx, out, nullptr, nullptr, in, n, n_ext, ncols);
}

// ------------------------------------------------------------------------------------------------
// Poseidon / Merkle
// ------------------------------------------------------------------------------------------------
extern "C" uint64_t glp_tree_num_elements(uint64_t nrows) { return nrows ? (2 * nrows - 1) * 4 : 0; }

static int dev_leaf_hash(glp_ctx *ctx, gl_t *out4, const gl_t *in, uint64_t ncols, uint64_t pitch, uint64_t nrows, const gl_t *cap_in = nullptr,
                         uint64_t out_pitch = 4)
{
    if (!nrows) return GLP_OK;
    const int vec16 = (pitch % 2 == 0) && (((uintptr_t)in) % 16 == 0);
    const uint64_t nblk = (nrows + POS_LEAF_BLOCK - 1) / POS_LEAF_BLOCK;
    if (nblk > 0x7FFFFFFFull) return fail(GLP_E_ARG, "too many rows");
    LAUNCH_FLAT(ctx, k_linear_hash_rows<POS_LEAF_LOCKSTEP != 0>, (unsigned)nblk, POS_LEAF_BLOCK, out4, in, ncols, pitch, nrows, vec16, cap_in, out_pitch);
    CU(cudaGetLastError());
    return GLP_OK;
}
static int dev_tree_levels(glp_ctx *ctx, gl_t *level, uint64_t n)
{
    gl_t *cur = level;
    while (n > 2 * POS_TOP_THREADS) {
        gl_t *par = cur + 4 * n;
        const uint64_t np = n / 2;
        LAUNCH_FLAT(ctx, k_tree_level, (unsigned)((np + POS_BLOCK - 1) / POS_BLOCK), POS_BLOCK, par, (const gl_t *)cur, np);
        cur = par;
        n = np;
    }
    if (n > 1) LAUNCH(ctx, k_tree_top, 1, POS_TOP_THREADS, 0, cur, n);
    CU(cudaGetLastError());
    return GLP_OK;
}
static int dev_merkletree(glp_ctx *ctx, gl_t *tree, const gl_t *in, uint64_t width, uint64_t pitch, uint64_t nrows)
{
    {
        ProfScope ps(ctx, 2);
        GLP(dev_leaf_hash(ctx, tree, in, width, pitch, nrows));
    }
    {
        ProfScope ps(ctx, 3);
        GLP(dev_tree_levels(ctx, tree, nrows));
    }
    return GLP_OK;
}

extern "C" int glp_hash_full_result(glp_ctx *ctx, uint64_t *out, const uint64_t *in, uint64_t n)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (!n) return GLP_OK;
    if (!out || !in) return fail(GLP_E_ARG, "glp_hash_full_result: NULL pointer");
    CU(cudaSetDevice(ctx->device));
    const gl_t *din = nullptr;
    GLP(stage_in(ctx, ctx->tmp_in, in, 12 * n, &din));
    const bool out_dev = on_device(out);
    gl_t *dout = (gl_t *)out;
    if (!out_dev) { GLP(ensure(ctx, ctx->tmp_out, 12 * n * sizeof(gl_t))); dout = (gl_t *)ctx->tmp_out.p; }
    LAUNCH_FLAT(ctx, k_permute_batch, (unsigned)((n + POS_BLOCK - 1) / POS_BLOCK), POS_BLOCK, dout, din, n);
    CU(cudaGetLastError());
    if (!out_dev) CU(cudaMemcpyAsync(out, dout, 12 * n * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

extern "C" int glp_linear_hash_rows(glp_ctx *ctx, uint64_t *out4, const uint64_t *in, uint64_t ncols, uint64_t pitch, uint64_t nrows)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (!nrows) return GLP_OK;
    if (!out4 || (!in && ncols)) return fail(GLP_E_ARG, "glp_linear_hash_rows: NULL pointer");
    if (pitch < ncols) return fail(GLP_E_ARG, "glp_linear_hash_rows: pitch smaller than ncols");
    CU(cudaSetDevice(ctx->device));
    const gl_t *din = (const gl_t *)in;
    if (ncols) GLP(stage_in(ctx, ctx->tmp_in, in, (size_t)(nrows - 1) * pitch + ncols, &din));
    const bool out_dev = on_device(out4);
    gl_t *dout = (gl_t *)out4;
    if (!out_dev) { GLP(ensure(ctx, ctx->tmp_out, 4 * nrows * sizeof(gl_t))); dout = (gl_t *)ctx->tmp_out.p; }
    GLP(dev_leaf_hash(ctx, dout, din, ncols, pitch, nrows));
    if (!out_dev) CU(cudaMemcpyAsync(out4, dout, 4 * nrows * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

extern "C" int glp_linear_hash_absorb(glp_ctx *ctx, uint64_t *state4, const uint64_t *in, uint64_t ncols, uint64_t pitch, uint64_t nrows,
                                      int first)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (!nrows) return GLP_OK;
    if (!state4 || (!in && ncols)) return fail(GLP_E_ARG, "glp_linear_hash_absorb: NULL pointer");
    if (pitch < ncols) return fail(GLP_E_ARG, "glp_linear_hash_absorb: pitch smaller than ncols");
    if (!first && ncols == 0) return GLP_OK;                        // nothing to absorb
    CU(cudaSetDevice(ctx->device));
    const gl_t *din = (const gl_t *)in;
    if (ncols) GLP(stage_in(ctx, ctx->tmp_in, in, (size_t)(nrows - 1) * pitch + ncols, &din));
    const bool st_dev = on_device(state4);
    gl_t *dst = (gl_t *)state4;
    if (!st_dev) {
        GLP(ensure(ctx, ctx->tmp_out, 4 * nrows * sizeof(gl_t)));
        dst = (gl_t *)ctx->tmp_out.p;
        if (!first) CU(cudaMemcpyAsync(dst, state4, 4 * nrows * sizeof(gl_t), cudaMemcpyHostToDevice, ctx->stream));
    }
    GLP(dev_leaf_hash(ctx, dst, din, ncols, pitch, nrows, first ? nullptr : dst));
    if (!st_dev) CU(cudaMemcpyAsync(state4, dst, 4 * nrows * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

// Host-resident source (what MerkleTreeGL::merkelize() hands over when starks.cpp is left unchanged): rows are independent
// leaves, so the matrix is uploaded in contiguous row blocks of ~1 GB on a copy stream while the previous block is hashed;
// the upload (52-57 GB/s) and the leaf hash (65 GB/s of input at 660 columns) overlap and two staging blocks replace a
// device copy of the whole matrix.
#define GLP_MERKLE_STREAM_MIN_BYTES ((uint64_t)256 << 20)
static int merkletree_streamed(glp_ctx *ctx, gl_t *dtree, const uint64_t *in, uint64_t width, uint64_t nrows)
{
    const uint64_t target = ctx->merkle_stream_min ? ctx->merkle_stream_min / 4 + 1 : ((uint64_t)1 << 30);   // bytes per block
    uint64_t rb = nrows;
    while (rb > 1024 && rb * width * sizeof(gl_t) > target) rb >>= 1;
    const uint64_t nblocks = nrows / rb;                                     // both powers of two
    const size_t block_words = (size_t)rb * width;
    GLP(ensure(ctx, ctx->tmp_in, 2 * block_words * sizeof(gl_t)));
    gl_t *stage[2] = {(gl_t *)ctx->tmp_in.p, (gl_t *)ctx->tmp_in.p + block_words};
    std::vector<cudaEvent_t> up(nblocks), hashed(nblocks);
    PipelineEvents pe(ctx);
    for (uint64_t k = 0; k < nblocks; k++) {
        GLP(pe.make(&up[k], cudaEventDisableTiming));
        GLP(pe.make(&hashed[k], cudaEventDisableTiming));
    }
    cudaEvent_t start;
    GLP(pe.make(&start, cudaEventDisableTiming));
    CU(cudaEventRecord(start, ctx->stream));                                  // the staging buffers may still be in use by earlier work
    CU(cudaStreamWaitEvent(ctx->copy_stream[0], start, 0));
    int rc = GLP_OK;
    {
        ProfScope ps(ctx, 2);
        for (uint64_t k = 0; k < nblocks && rc == GLP_OK; k++) {
            if (k >= 2) CU(cudaStreamWaitEvent(ctx->copy_stream[0], hashed[k - 2], 0));
            CU(cudaMemcpyAsync(stage[k & 1], in + k * block_words, block_words * sizeof(gl_t), cudaMemcpyHostToDevice, ctx->copy_stream[0]));
            CU(cudaEventRecord(up[k], ctx->copy_stream[0]));
            CU(cudaStreamWaitEvent(ctx->stream, up[k], 0));
            rc = dev_leaf_hash(ctx, dtree + 4 * k * rb, stage[k & 1], width, width, rb);
            CU(cudaEventRecord(hashed[k], ctx->stream));
        }
    }
    if (rc == GLP_OK) {
        ProfScope ps(ctx, 3);
        rc = dev_tree_levels(ctx, dtree, nrows);
    }
    CU(cudaStreamSynchronize(ctx->copy_stream[0]));
    return rc;
}

extern "C" int glp_merkletree(glp_ctx *ctx, uint64_t *tree, const uint64_t *in, uint64_t ncols, uint64_t nrows, uint64_t dim)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (nrows == 0) return GLP_OK;                                  // poseidon_goldilocks.cpp:268-271
    const uint64_t width = ncols * dim;
    if (!tree || (!in && width)) return fail(GLP_E_ARG, "glp_merkletree: NULL pointer");
    if (!is_pow2(nrows)) return fail(GLP_E_ARG, "glp_merkletree: nrows must be a power of two");
    CU(cudaSetDevice(ctx->device));
    const size_t tw = glp_tree_num_elements(nrows);
    const bool out_dev = on_device(tree);
    gl_t *dtree = (gl_t *)tree;
    if (!out_dev) { GLP(ensure(ctx, ctx->nodes, tw * sizeof(gl_t))); dtree = (gl_t *)ctx->nodes.p; }
    const uint64_t stream_min = ctx->merkle_stream_min ? ctx->merkle_stream_min : GLP_MERKLE_STREAM_MIN_BYTES;
    if (width > POS_CAP && !on_device(in) && nrows >= 2048 && nrows * width * sizeof(gl_t) >= stream_min) {
        GLP(merkletree_streamed(ctx, dtree, in, width, nrows));
    } else {
        const gl_t *din = (const gl_t *)in;
        if (width) GLP(stage_in(ctx, ctx->tmp_in, in, (size_t)nrows * width, &din));
        GLP(dev_merkletree(ctx, dtree, din, width, width, nrows));
    }
    if (!out_dev) CU(cudaMemcpyAsync(tree, dtree, tw * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

extern "C" int glp_merkletree_batch(glp_ctx *ctx, uint64_t *tree, const uint64_t *in, uint64_t ncols, uint64_t nrows, uint64_t batch_size,
                                    uint64_t dim)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (nrows == 0) return GLP_OK;                                  // poseidon_goldilocks.cpp:115-118
    const uint64_t width = ncols * dim;
    if (!tree || (!in && width)) return fail(GLP_E_ARG, "glp_merkletree_batch: NULL pointer");
    if (!is_pow2(nrows)) return fail(GLP_E_ARG, "glp_merkletree_batch: nrows must be a power of two");
    if (batch_size == 0) return fail(GLP_E_ARG, "glp_merkletree_batch: batch_size is 0");
    CU(cudaSetDevice(ctx->device));
    const uint64_t nb = ncols ? (ncols + batch_size - 1) / batch_size : 1;      // :121-126
    const uint64_t nlast = ncols - (nb - 1) * batch_size;
    const size_t tw = glp_tree_num_elements(nrows);
    const gl_t *din = (const gl_t *)in;
    if (width) GLP(stage_in(ctx, ctx->tmp_in, in, (size_t)nrows * width, &din));
    const bool out_dev = on_device(tree);
    gl_t *dtree = (gl_t *)tree;
    if (!out_dev) { GLP(ensure(ctx, ctx->nodes, tw * sizeof(gl_t))); dtree = (gl_t *)ctx->nodes.p; }
    // per-batch hashes side by side: [nrows x 4*nb], then one more linear_hash over each of those rows (:131-141)
    GLP(ensure(ctx, ctx->tmp_out, (size_t)nrows * 4 * nb * sizeof(gl_t)));
    gl_t *bh = (gl_t *)ctx->tmp_out.p;
    {
        ProfScope ps(ctx, 2);
        for (uint64_t j = 0; j < nb; j++)
            GLP(dev_leaf_hash(ctx, bh + 4 * j, din + j * batch_size * dim, (j == nb - 1 ? nlast : batch_size) * dim, width, nrows, nullptr, 4 * nb));
        GLP(dev_leaf_hash(ctx, dtree, bh, 4 * nb, 4 * nb, nrows));
    }
    {
        ProfScope ps(ctx, 3);
        GLP(dev_tree_levels(ctx, dtree, nrows));
    }
    if (!out_dev) CU(cudaMemcpyAsync(tree, dtree, tw * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

extern "C" int glp_merkle_levels(glp_ctx *ctx, uint64_t *level, uint64_t n)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (n <= 1) return GLP_OK;
    if (!level) return fail(GLP_E_ARG, "glp_merkle_levels: NULL pointer");
    if (!is_pow2(n)) return fail(GLP_E_ARG, "glp_merkle_levels: n must be a power of two");
    CU(cudaSetDevice(ctx->device));
    const size_t words = (size_t)(2 * n - 1) * 4;
    const bool dev = on_device(level);
    gl_t *d = (gl_t *)level;
    if (!dev) {
        GLP(ensure(ctx, ctx->nodes, words * sizeof(gl_t)));
        d = (gl_t *)ctx->nodes.p;
        CU(cudaMemcpyAsync(d, level, 4 * n * sizeof(gl_t), cudaMemcpyHostToDevice, ctx->stream));
    }
    GLP(dev_tree_levels(ctx, d, n));
    if (!dev) CU(cudaMemcpyAsync(level + 4 * n, d + 4 * n, (words - 4 * n) * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

extern "C" int glp_merkle_group_proof(glp_ctx *ctx, uint64_t *proof, const uint64_t *source, const uint64_t *nodes, uint64_t width,
                                      uint64_t height, uint64_t idx)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (!proof || !nodes || (!source && width)) return fail(GLP_E_ARG, "glp_merkle_group_proof: NULL pointer");
    if (!is_pow2(height) || idx >= height) return fail(GLP_E_ARG, "glp_merkle_group_proof: bad height/idx");
    const uint64_t depth = ilog2_u64(height);
    const uint64_t words = width + 4 * depth;
    if (!on_device(source) || !on_device(nodes)) {
        // host-resident tree: this is a pure gather, do it where the data is (merkleTreeGL.cpp:12-35)
        for (uint64_t i = 0; i < width; i++) proof[i] = gl_canon(source[idx * width + i]);
        uint64_t n = height, off = 0, id = idx;
        for (uint64_t l = 0; l < depth; l++) {
            memcpy(proof + width + 4 * l, nodes + off + (id ^ 1) * 4, 4 * sizeof(uint64_t));
            off += 4 * n; n >>= 1; id >>= 1;
        }
        return GLP_OK;
    }
    CU(cudaSetDevice(ctx->device));
    GLP(ensure(ctx, ctx->tmp_out, words * sizeof(gl_t)));
    LAUNCH(ctx, k_group_proof, 1, 128, 0, (gl_t *)ctx->tmp_out.p, (const gl_t *)source, (const gl_t *)nodes, width, height, idx);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(proof, ctx->tmp_out.p, words * sizeof(gl_t), cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

// All the proofs one tree owes to the FRI query phase in ONE launch and ONE download (friProve.cpp:150-170 asks each of the
// five trees for every query index through queryPol, :219-250): proofs[q * (width + 4*depth) ...] = getGroupProof(idx[q]).
// With source and nodes resident on the device this is what makes "keep cm1..cm4 on the device, download only roots,
// query rows and paths" usable: nidx * (width + 4*depth) words cross the link instead of the whole extension.
extern "C" int glp_merkle_group_proofs(glp_ctx *ctx, uint64_t *proofs, const uint64_t *source, const uint64_t *nodes, uint64_t width,
                                       uint64_t height, const uint64_t *idx, uint64_t nidx)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (nidx == 0) return GLP_OK;
    if (!proofs || !nodes || !idx || (!source && width)) return fail(GLP_E_ARG, "glp_merkle_group_proofs: NULL pointer");
    if (!is_pow2(height)) return fail(GLP_E_ARG, "glp_merkle_group_proofs: height must be a power of two");
    if (on_device(idx)) return fail(GLP_E_ARG, "glp_merkle_group_proofs: idx is a host array");
    for (uint64_t q = 0; q < nidx; q++)
        if (idx[q] >= height) return fail(GLP_E_ARG, "glp_merkle_group_proofs: index out of range");
    const uint64_t depth = ilog2_u64(height);
    const uint64_t words = width + 4 * depth;
    if (!on_device(source) || !on_device(nodes)) {
        if (on_device(source) || on_device(nodes) || on_device(proofs))
            return fail(GLP_E_ARG, "glp_merkle_group_proofs: source and nodes must both be host or both be device memory");
        // host-resident tree: a pure gather, done where the data is (merkleTreeGL.cpp:12-35)
        for (uint64_t q = 0; q < nidx; q++) {
            uint64_t *proof = proofs + q * words;
            for (uint64_t i = 0; i < width; i++) proof[i] = gl_canon(source[idx[q] * width + i]);
            uint64_t n = height, off = 0, id = idx[q];
            for (uint64_t l = 0; l < depth; l++) {
                memcpy(proof + width + 4 * l, nodes + off + (id ^ 1) * 4, 4 * sizeof(uint64_t));
                off += 4 * n; n >>= 1; id >>= 1;
            }
        }
        return GLP_OK;
    }
    CU(cudaSetDevice(ctx->device));
    if (nidx > 0x7FFFFFFFull) return fail(GLP_E_ARG, "glp_merkle_group_proofs: too many indices");
    const bool out_dev = on_device(proofs);
    const size_t out_words = (size_t)nidx * words;
    // staging: [indices | proofs]
    GLP(ensure(ctx, ctx->tmp_out, (nidx + (out_dev ? 0 : out_words)) * sizeof(gl_t)));
    uint64_t *didx = (uint64_t *)ctx->tmp_out.p;
    gl_t *dproofs = out_dev ? (gl_t *)proofs : (gl_t *)ctx->tmp_out.p + nidx;
    CU(cudaMemcpyAsync(didx, idx, nidx * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
    LAUNCH(ctx, k_group_proofs, (unsigned)nidx, 128, 0, dproofs, words, (const gl_t *)source, (const gl_t *)nodes, width, height,
           (const uint64_t *)didx, nidx);
    CU(cudaGetLastError());
    if (!out_dev) CU(cudaMemcpyAsync(proofs, dproofs, out_words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

// ------------------------------------------------------------------------------------------------
// fused commit
// ------------------------------------------------------------------------------------------------
// Column-block boundaries of the streamed commit: col0[0] = 0 < ... < col0[nblocks] = ncols, every boundary but the last a
// multiple of 8 (whole sponge blocks), no block wider than wmax.
//   * GLP_STREAM_PLAN="32,48,72,..." gives the widths (the last one repeats until the columns are used up);
//   * GLP_STREAM_GROW=0 is the round-1 layout: one short first block (32 columns), then equal blocks;
//   * default (GLP_STREAM_GROW=200): a graded start -- 32, 64, 128 columns, each block twice the one before until wmax is
//     reached or half of the rest is used -- then equal blocks (660 columns: 32, 64, 128, 152, 144, 140).  The download is the
//     longest stream of the pipeline (2 x the upload's bytes over the same link) and block k's download can only start when
//     block k's LDE is done, which waits for block k's upload: with equal 160-column blocks the download link idles for most
//     of the first 0.35 s.  Measured at 2^23 x 660 on one box (profiles/r02_e2e_plan.txt): 1983 ms equal blocks, 1934 ms
//     graded x2, 1928 ms for 32,64,96,160,160,148, 1982-1997 ms for finer starts (x1.5: 32,48,72,112,...) whose narrow
//     blocks cost LDE lanes and short copy rows.  What is left is the link: 133 GB both ways at 82-86 GB/s (2-D + 2-D) while
//     the upload runs, 49.5 GB/s for the download alone afterwards (profiles/r01_pcie_probe.txt).
//     A short LAST block would only be copied in short, slow rows (160-byte rows move at 18 GB/s, 768-byte rows at 46-52 GB/s),
//     so the tail stays equal.
static std::vector<uint64_t> stream_block_plan(const glp_ctx *ctx, uint64_t ncols, uint64_t wmax)
{
    const uint64_t units = (ncols + 7) / 8, wu = wmax / 8 ? wmax / 8 : 1;
    std::vector<uint64_t> u;                                                 // block sizes in units of 8 columns
    uint64_t rem = units;
    if (!ctx->stream_plan.empty()) {
        for (size_t i = 0; rem; i++) {
            uint64_t x = ctx->stream_plan[i < ctx->stream_plan.size() ? i : ctx->stream_plan.size() - 1] / 8;
            x = x < 1 ? 1 : (x > wu ? wu : x);
            x = x > rem ? rem : x;
            u.push_back(x);
            rem -= x;
        }
    } else {
        if (ctx->stream_grow == 0) {
            if (units > 2 * wu && wu > 4) { u.push_back(4); rem -= 4; }
        } else {
            for (uint64_t x = 4; x < wu && rem >= 2 * x; x = (x * ctx->stream_grow + 99) / 100) { u.push_back(x); rem -= x; }
        }
        const uint64_t nrest = rem ? (rem + wu - 1) / wu : 0;
        for (uint64_t k = 0; k < nrest; k++) u.push_back(rem / nrest + (k < rem % nrest ? 1 : 0));
    }
    std::vector<uint64_t> col0(u.size() + 1);
    col0[0] = 0;
    for (size_t k = 0; k < u.size(); k++) col0[k + 1] = (col0[k] + 8 * u[k] < ncols) ? col0[k] + 8 * u[k] : ncols;
    col0.back() = ncols;
    return col0;
}

// Host input: the commit as a software pipeline over column blocks of `w` columns (a multiple of 8, so every
// sponge block of the leaf hash has one owner).  Per block k:
//     copy stream 0 : upload block k (strided 2-D copy out of the row-major host matrix)
//     compute stream: LDE of block k  ->  absorb block k into every row's sponge (the leaf array IS the sponge state)
//     copy stream 1 : download block k of the extension (strided 2-D copy into the row-major host matrix)
// so H2D, arithmetic and D2H of different blocks overlap and the full extended matrix never has to exist on the
// device (2 x (N + N_ext) x w words of staging instead).  After the last block the leaf array holds the leaf
// hashes and the level loop runs in place.  out_ext may alias in (same pitch): block k's download only overwrites
// the columns block k's upload has already read.
static int commit_streamed(glp_ctx *ctx, uint64_t *out_ext, bool ext_dev, uint64_t *tree, bool tree_dev, gl_t *dtree, uint64_t root[4],
                           const uint64_t *in, uint64_t n, uint64_t n_ext, uint64_t ncols, uint64_t in_pitch = 0, uint64_t out_pitch = 0)
{
    if (!in_pitch) in_pitch = ncols;                              // a column window of a wider host matrix (glp_multi_extend_pol)
    if (!out_pitch) out_pitch = ncols;
    const bool want_tree = dtree != nullptr;
    const bool want_ext = out_ext != nullptr;
    const size_t tree_words = glp_tree_num_elements(n_ext);
    const uint32_t logN = ilog2_u64(n), logE = ilog2_u64(n_ext);
    // block width: GLP_STREAM_COLS (default 160), shrunk to fit the scratch budget
    uint64_t w = ctx->stream_cols;
    bool full_stage = false;
    {
        size_t f = 0, t = 0;
        if (cudaMemGetInfo(&f, &t) != cudaSuccess) { cudaGetLastError(); f = (size_t)1 << 30; }
        uint64_t budget = (uint64_t)f + ctx->tmp_in.bytes + ctx->tmp_out.bytes + ctx->s2.bytes;
        budget = budget / 10 * 8;
        if (ctx->scratch_limit && ctx->scratch_limit < budget) budget = ctx->scratch_limit;
        const uint64_t per_col = 8 * (2 * n + (ext_dev ? 0 : 2 * n_ext) + n);
        while (w > 8 && w * per_col > budget) w -= 8;
        // When the whole input also fits, every block gets its own staging slice and the upload runs ahead at full
        // rate instead of in step with the LDE: the download -- the longest stream of the pipeline -- then shares the
        // PCIe link with an upload for the first 0.85 s only (2-D downloads: 43 GB/s next to an upload, 47 GB/s alone).
        full_stage = ctx->stage_all && 8 * n * (ncols + 8) + (w + 8) * (per_col - 16 * n) <= budget;
    }
    const std::vector<uint64_t> col0 = stream_block_plan(ctx, ncols, w);
    const uint64_t nblocks = col0.size() - 1;
    w = 0;                                                                   // widest block: sizes the buffers
    for (uint64_t k = 0; k < nblocks; k++) w = (col0[k + 1] - col0[k] > w) ? col0[k + 1] - col0[k] : w;
    w = (w + 7) / 8 * 8;
    GLP(ensure(ctx, ctx->tmp_in, (full_stage ? (size_t)n * (col0[nblocks] + 8) : 2 * (size_t)n * w) * sizeof(gl_t)));
    if (!ext_dev) GLP(ensure(ctx, ctx->tmp_out, 2 * (size_t)n_ext * w * sizeof(gl_t)));
    // the LDE scratch for the WIDEST block, before the streams fork: lde_block's own ensure() would otherwise free and
    // reallocate it between the short first block and the next one -- cudaFree synchronises the whole device, uploads included
    GLP(ensure(ctx, ctx->s2, (size_t)n * w * sizeof(gl_t)));
    if (ctx->no_fused_lde) GLP(ensure(ctx, ctx->s1, (size_t)n_ext * w * sizeof(gl_t)));
    auto stage_of = [&](uint64_t k) { return (gl_t *)ctx->tmp_in.p + (full_stage ? (size_t)n * col0[k] : (k & 1) * (size_t)n * w); };
    gl_t *extb[2] = {(gl_t *)ctx->tmp_out.p, ext_dev ? nullptr : (gl_t *)ctx->tmp_out.p + (size_t)n_ext * w};
    std::vector<cudaEvent_t> up(nblocks), lde_done(nblocks), down(nblocks), hashed(nblocks), half(nblocks);
    const unsigned evflags = ctx->trace ? cudaEventDefault : cudaEventDisableTiming;     // GLP_TRACE=1: print the pipeline's timeline
    PipelineEvents pe(ctx);
    for (uint64_t k = 0; k < nblocks; k++) {
        GLP(pe.make(&up[k], evflags));
        GLP(pe.make(&lde_done[k], evflags));
        GLP(pe.make(&down[k], evflags));
        GLP(pe.make(&hashed[k], evflags));
        GLP(pe.make(&half[k], cudaEventDisableTiming));
    }
    cudaEvent_t start;
    GLP(pe.make(&start, evflags));
    CU(cudaEventRecord(start, ctx->stream));                      // the copy streams must not run ahead of earlier work on the caller's stream
    CU(cudaStreamWaitEvent(ctx->copy_stream[0], start, 0));
    CU(cudaStreamWaitEvent(ctx->copy_stream[1], start, 0));
    CU(cudaStreamWaitEvent(ctx->copy_stream[2], start, 0));
    auto width = [&](uint64_t k) { return (uint32_t)(col0[k + 1] - col0[k]); };
    auto upload = [&](uint64_t k) -> int {
        if (k >= 2 && !full_stage) CU(cudaStreamWaitEvent(ctx->copy_stream[0], lde_done[k - 2], 0));   // staging buffer read by block k-2's INTT
        CU(cudaMemcpy2DAsync(stage_of(k), (size_t)width(k) * 8, in + col0[k], (size_t)in_pitch * 8, (size_t)width(k) * 8, n, cudaMemcpyHostToDevice,
                             ctx->copy_stream[0]));
        CU(cudaEventRecord(up[k], ctx->copy_stream[0]));
        return GLP_OK;
    };
    GLP(upload(0));
    if (nblocks > 1) GLP(upload(1));
    if (full_stage)
        for (uint64_t k = 2; k < nblocks; k++) GLP(upload(k));
    for (uint64_t k = 0; k < nblocks; k++) {
        const uint64_t c0 = col0[k];
        const uint32_t wk = width(k);
        gl_t *eb = ext_dev ? (gl_t *)out_ext + c0 : extb[k & 1];
        const uint64_t eb_pitch = ext_dev ? out_pitch : wk;
        CU(cudaStreamWaitEvent(ctx->stream, up[k], 0));
        if (!ext_dev && want_ext && k >= 2) CU(cudaStreamWaitEvent(ctx->stream, down[k - 2], 0));   // extension buffer still being downloaded
        GLP(lde_block(ctx, logN, logE, stage_of(k), wk, eb, eb_pitch, wk));
        CU(cudaEventRecord(lde_done[k], ctx->stream));
        if (k + 2 < nblocks && !full_stage) GLP(upload(k + 2));
        if (want_ext && !ext_dev) {
            // one strided 2-D copy keeps a single copy engine busy at 45-49 GB/s; two halves of the rows on two streams run on
            // two engines (GLP_D2H_SPLIT, profiles/r02_e2e_trace.txt)
            const uint64_t r1 = (ctx->d2h_split == 2 && n_ext >= 4096) ? n_ext / 2 : n_ext;
            CU(cudaStreamWaitEvent(ctx->copy_stream[1], lde_done[k], 0));
            CU(cudaMemcpy2DAsync(out_ext + c0, (size_t)out_pitch * 8, eb, (size_t)wk * 8, (size_t)wk * 8, r1, cudaMemcpyDeviceToHost,
                                 ctx->copy_stream[1]));
            if (r1 < n_ext) {
                CU(cudaStreamWaitEvent(ctx->copy_stream[2], lde_done[k], 0));
                CU(cudaMemcpy2DAsync(out_ext + r1 * out_pitch + c0, (size_t)out_pitch * 8, eb + r1 * eb_pitch, (size_t)wk * 8, (size_t)wk * 8,
                                     n_ext - r1, cudaMemcpyDeviceToHost, ctx->copy_stream[2]));
                CU(cudaEventRecord(half[k], ctx->copy_stream[2]));
                CU(cudaStreamWaitEvent(ctx->copy_stream[1], half[k], 0));   // down[k] = both halves
            }
            CU(cudaEventRecord(down[k], ctx->copy_stream[1]));
        }
        if (want_tree) {
            ProfScope ps(ctx, 2);
            GLP(dev_leaf_hash(ctx, dtree, eb, wk, eb_pitch, n_ext, k == 0 ? nullptr : dtree));
        }
        if (ctx->trace) CU(cudaEventRecord(hashed[k], ctx->stream));
    }
    if (want_tree) {
        {
            ProfScope ps(ctx, 3);
            GLP(dev_tree_levels(ctx, dtree, n_ext));
        }
        if (tree && !tree_dev) CU(cudaMemcpyAsync(tree, dtree, tree_words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
        if (root) CU(cudaMemcpyAsync(root, dtree + tree_words - 4, 4 * sizeof(gl_t), cudaMemcpyDefault, ctx->stream));
    }
    cudaError_t e0 = cudaStreamSynchronize(ctx->stream), e1 = cudaStreamSynchronize(ctx->copy_stream[0]),
                e2 = cudaStreamSynchronize(ctx->copy_stream[1]);
    if (ctx->trace && e0 == cudaSuccess && e1 == cudaSuccess && e2 == cudaSuccess) {
        fprintf(stderr, "[glp trace] streamed commit: %llu blocks of at most %llu columns; ms since start: block  uploaded  extended  hashed  downloaded\n",
                (unsigned long long)nblocks, (unsigned long long)w);
        for (uint64_t k = 0; k < nblocks; k++) {
            float t[4] = {-1, -1, -1, -1};
            cudaEventElapsedTime(&t[0], start, up[k]);
            cudaEventElapsedTime(&t[1], start, lde_done[k]);
            if (want_tree) cudaEventElapsedTime(&t[2], start, hashed[k]);
            if (want_ext && !ext_dev) cudaEventElapsedTime(&t[3], start, down[k]);
            fprintf(stderr, "[glp trace] %3llu %9.1f %9.1f %9.1f %9.1f   (%u columns)\n", (unsigned long long)k, t[0], t[1], t[2], t[3], width(k));
        }
        cudaGetLastError();
    }
    CU(e0); CU(e1); CU(e2);
    return GLP_OK;
}

// Device-resident commit with the two stages overlapped ON THE SAME SMs.  The leaf hash is bound by the multiplier pipes
// (IMAD.WIDE / DFMA) and leaves ~40 % of the issue slots and of the ALU pipe idle; the NTT passes are bound by the ALU
// pipe and use the multiplier lightly.  So the matrix is cut into column blocks (multiples of 8 columns:
// whole sponge blocks), the main stream absorbs block k into every row's sponge (the leaf array is the sponge state,
// glp_linear_hash_absorb) while a high-priority stream extends block k+1 with k_ntt_pass_co -- a small grid that is
// resident next to the hash CTAs.  Block 0 is short and extended with the full-size kernels (nothing to hide it under).
static int commit_overlapped(glp_ctx *ctx, gl_t *dext, gl_t *dtree, const gl_t *in, uint64_t n, uint64_t n_ext, uint64_t ncols)
{
    const uint32_t logN = ilog2_u64(n), logE = ilog2_u64(n_ext);
    std::vector<uint64_t> col0;
    {
        uint64_t c = 0;
        col0.push_back(0);
        if (ctx->co_first_cols && ncols > 2 * (uint64_t)ctx->co_first_cols) { c = ctx->co_first_cols; col0.push_back(c); }
        const uint64_t rest = ncols - c, wmax = ctx->co_cols;
        const uint64_t nb = (rest + wmax - 1) / wmax;
        const uint64_t units = (rest + 7) / 8;                               // equal blocks, in units of 8 columns
        for (uint64_t k = 0; k < nb; k++) {
            const uint64_t u = units / nb + (k < units % nb ? 1 : 0);
            c += u * 8;
            col0.push_back(c < ncols ? c : ncols);
        }
        col0.back() = ncols;
    }
    const uint64_t nblocks = col0.size() - 1;
    uint64_t wmaxb = 0;
    for (uint64_t k = 0; k < nblocks; k++) wmaxb = (col0[k + 1] - col0[k] > wmaxb) ? col0[k + 1] - col0[k] : wmaxb;
    GLP(ensure(ctx, ctx->s2, (size_t)n * wmaxb * sizeof(gl_t)));              // before the streams fork: no allocation inside the pipeline
    cudaStream_t main_stream = ctx->stream;
    cudaEvent_t start, done;
    std::vector<cudaEvent_t> ext_done(nblocks);
    CU(cudaEventCreateWithFlags(&start, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
    for (auto &e : ext_done) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CU(cudaEventRecord(start, main_stream));
    CU(cudaStreamWaitEvent(ctx->lde_stream, start, 0));
    int rc = GLP_OK;
    for (uint64_t k = 0; k < nblocks && rc == GLP_OK; k++) {
        const uint64_t c0 = col0[k];
        const uint32_t wk = (uint32_t)(col0[k + 1] - c0);
        ctx->stream = ctx->lde_stream;                                         // the launches of lde_block go to the high-priority stream
        ctx->co_active = (k > 0 && ctx->co_mode == 1) ? 1 : 0;
        rc = lde_block(ctx, logN, logE, in + c0, ncols, dext + c0, ncols, wk);
        ctx->co_active = 0;
        ctx->stream = main_stream;
        if (rc != GLP_OK) break;
        CU(cudaEventRecord(ext_done[k], ctx->lde_stream));
        CU(cudaStreamWaitEvent(main_stream, ext_done[k], 0));
        ProfScope ps(ctx, 2);
        rc = dev_leaf_hash(ctx, dtree, dext + c0, wk, ncols, n_ext, k == 0 ? nullptr : dtree);
    }
    if (rc == GLP_OK) {
        ProfScope ps(ctx, 3);
        rc = dev_tree_levels(ctx, dtree, n_ext);
    }
    // the caller's stream also covers whatever is left on the LDE stream (nothing, unless a launch failed half way)
    cudaEventRecord(done, ctx->lde_stream);
    cudaStreamWaitEvent(main_stream, done, 0);
    cudaEventDestroy(start);
    cudaEventDestroy(done);
    for (auto &e : ext_done) cudaEventDestroy(e);
    return rc;
}

extern "C" int glp_commit(glp_ctx *ctx, uint64_t *out_ext, uint64_t *tree, uint64_t root[4], const uint64_t *in, uint64_t n,
                          uint64_t n_ext, uint64_t ncols)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (n == 0 || ncols == 0) return GLP_OK;
    if (!in) return fail(GLP_E_ARG, "glp_commit: in is NULL");
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GLP_E_ARG, "glp_commit: sizes must be powers of two with n_ext >= n");
    if (n_ext > (1ull << 32)) return fail(GLP_E_DOMAIN, "Domain size too big for the curve");
    CU(cudaSetDevice(ctx->device));
    const bool want_tree = tree != nullptr || root != nullptr;
    const size_t in_words = (size_t)n * ncols, ext_words = (size_t)n_ext * ncols, tree_words = glp_tree_num_elements(n_ext);
    const bool in_dev = on_device(in), ext_dev = on_device(out_ext), tree_dev = on_device(tree);

    gl_t *dtree = (gl_t *)tree;
    if (want_tree && !tree_dev) { GLP(ensure(ctx, ctx->nodes, tree_words * sizeof(gl_t))); dtree = (gl_t *)ctx->nodes.p; }
    if (!in_dev) return commit_streamed(ctx, out_ext, ext_dev, tree, tree_dev, dtree, root, in, n, n_ext, ncols);

    // device-resident input: the extended matrix lives on the device for the Merkle stage
    gl_t *dext = (gl_t *)out_ext;
    if (!ext_dev) { GLP(ensure(ctx, ctx->ext, ext_words * sizeof(gl_t))); dext = (gl_t *)ctx->ext.p; }
    (void)in_words;
    // overlap the stages when the matrix is wide enough for a pipeline of column blocks and the LDE has a fused plan
    const bool overlap = want_tree && ctx->co_mode != 0 && !ctx->no_fused_lde && ncols >= 2 * (uint64_t)ctx->co_cols && dext != (const gl_t *)in;
    if (overlap) {
        GLP(commit_overlapped(ctx, dext, dtree, (const gl_t *)in, n, n_ext, ncols));
        if (out_ext && !ext_dev) CU(cudaMemcpyAsync(out_ext, dext, ext_words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    } else {
        // Optimistic LDE (with_optimistic_ntt), but with the flag read at the END of the commit, behind the one synchronise
        // the call makes anyway: the hash of an extension that turns out to be flagged is wasted, which never happens on
        // honest data, and the common path has no host round trip between the LDE and the hash.
        const gl_t *din = (const gl_t *)in, *in_end = din + (n - 1) * ncols + ncols, *ext_end = dext + ext_words;
        bool opt = ctx->opt_enabled && !ctx->no_fused_lde && (in_end <= dext || ext_end <= din);
        for (;;) {
            if (opt) {
                if (!ctx->opt_flag_dev) {
                    CU(cudaMalloc((void **)&ctx->opt_flag_dev, sizeof(uint32_t)));
                    CU(cudaHostAlloc((void **)&ctx->opt_flag_host, sizeof(uint32_t), cudaHostAllocDefault));
                }
                CU(cudaMemsetAsync(ctx->opt_flag_dev, 0, sizeof(uint32_t), ctx->stream));
            }
            ctx->opt_active = opt;
            const int rc = dev_extend_pol(ctx, dext, ncols, din, ncols, n_ext, n, ncols);
            ctx->opt_active = false;
            if (rc != GLP_OK) return rc;
            if (opt) CU(cudaMemcpyAsync(ctx->opt_flag_host, ctx->opt_flag_dev, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
            if (out_ext && !ext_dev) CU(cudaMemcpyAsync(out_ext, dext, ext_words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
            if (want_tree) GLP(dev_merkletree(ctx, dtree, dext, ncols, ncols, n_ext));
            if (tree && !tree_dev) CU(cudaMemcpyAsync(tree, dtree, tree_words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
            if (root) CU(cudaMemcpyAsync(root, dtree + tree_words - 4, 4 * sizeof(gl_t), cudaMemcpyDefault, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            if (!opt || !(*ctx->opt_flag_host || ctx->opt_force_retry)) return GLP_OK;
            ctx->opt_retries++;                                  // flagged: everything again with the exact passes
            opt = false;
        }
    }
    if (tree && !tree_dev) CU(cudaMemcpyAsync(tree, dtree, tree_words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    if (root) CU(cudaMemcpyAsync(root, dtree + tree_words - 4, 4 * sizeof(gl_t), cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

// ------------------------------------------------------------------------------------------------
// the callers either side of the commit
// ------------------------------------------------------------------------------------------------
// Step 4 of Starks::genProof (src/starkpil/starks.cpp:261-291): INTT of the quotient over the extended domain (qdim
// columns), chunk p of n coefficients scaled by shift^(-n p) into columns [p*qdim, (p+1)*qdim) (k_quotient_split),
// forward NTT of that n-row matrix zero-padded to n_ext rows, Merkle tree over the n_ext rows of the result.
// Everything between the input and cm4/tree stays on the device; narrow matrices (3 -> 6 columns) take the NARROW
// lane mapping of k_ntt_pass.
extern "C" int glp_quotient_commit(glp_ctx *ctx, uint64_t *cm4, uint64_t *tree, uint64_t root[4], const uint64_t *q, uint64_t n,
                                   uint64_t n_ext, uint64_t qdim, uint64_t qdeg)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (n == 0 || qdim == 0 || qdeg == 0) return GLP_OK;
    if (!q) return fail(GLP_E_ARG, "glp_quotient_commit: q is NULL");
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n || qdeg * n > n_ext)
        return fail(GLP_E_ARG, "glp_quotient_commit: sizes must be powers of two with qdeg * n <= n_ext");
    if (n_ext > (1ull << 32)) return fail(GLP_E_DOMAIN, "Domain size too big for the curve");
    if (qdim * qdeg > 0xFFFFu) return fail(GLP_E_ARG, "glp_quotient_commit: too many columns");
    CU(cudaSetDevice(ctx->device));
    const uint32_t logE = ilog2_u64(n_ext);
    const uint64_t w = qdim * qdeg;
    const size_t tree_words = glp_tree_num_elements(n_ext);
    const bool want_tree = tree != nullptr || root != nullptr;
    const bool cm_dev = on_device(cm4), tree_dev = on_device(tree);
    const gl_t *dq = nullptr;
    GLP(stage_in(ctx, ctx->tmp_in, q, (size_t)n_ext * qdim, &dq));
    GLP(ensure(ctx, ctx->work, ((size_t)n_ext * qdim + (size_t)n * w) * sizeof(gl_t)));
    gl_t *qq1 = (gl_t *)ctx->work.p, *qq2 = qq1 + (size_t)n_ext * qdim;
    GLP(ensure(ctx, ctx->s1, (size_t)n_ext * w * sizeof(gl_t)));
    gl_t *dcm = (gl_t *)cm4;
    if (!cm_dev) { GLP(ensure(ctx, ctx->ext, (size_t)n_ext * w * sizeof(gl_t))); dcm = (gl_t *)ctx->ext.p; }
    gl_t *dtree = (gl_t *)tree;
    if (want_tree && !tree_dev) { GLP(ensure(ctx, ctx->nodes, tree_words * sizeof(gl_t))); dtree = (gl_t *)ctx->nodes.p; }
    {
        ProfScope ps(ctx, 0);
        GLP(run_transform(ctx, logE, dq, qdim, qq1, qdim, (uint32_t)qdim, n_ext, true, false, (gl_t *)ctx->s1.p));
        const gl_t shift_in = gl_canon(gl_pow(gl_inv((gl_t)GL_SHIFT), n));
        const uint64_t nb = (n * w + 255) / 256;
        LAUNCH_FLAT(ctx, k_quotient_split, (unsigned)(nb > 148 * 32 ? 148 * 32 : nb), 256, qq2, (const gl_t *)qq1, n, (uint32_t)qdim, (uint32_t)qdeg, shift_in);
        CU(cudaGetLastError());
    }
    {
        ProfScope ps(ctx, 1);
        GLP(run_transform(ctx, logE, qq2, w, dcm, w, (uint32_t)w, n, false, false, (gl_t *)ctx->s1.p));
    }
    if (cm4 && !cm_dev) CU(cudaMemcpyAsync(cm4, dcm, (size_t)n_ext * w * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    if (want_tree) GLP(dev_merkletree(ctx, dtree, dcm, w, w, n_ext));
    if (tree && !tree_dev) CU(cudaMemcpyAsync(tree, dtree, tree_words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    if (root) CU(cudaMemcpyAsync(root, dtree + tree_words - 4, 4 * sizeof(gl_t), cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

// The constant tree in the layout bctree writes and MerkleTreeGL(Goldilocks::Element *tree) maps
// (tools/starkpil/bctree/build_const_tree.cpp:555-575; merkleTreeGL.hpp:24-32):
//     [npols, n_ext, extended constant polynomials (n_ext x npols), Merkle nodes ((2 n_ext - 1) x 4)]
// bctree's `interpolate` (:398-470) is the same map as extendPol, so this is glp_commit writing into one buffer.
extern "C" uint64_t glp_const_tree_num_elements(uint64_t n_ext, uint64_t npols) { return 2 + npols * n_ext + glp_tree_num_elements(n_ext); }
extern "C" int glp_const_tree(glp_ctx *ctx, uint64_t *const_tree, const uint64_t *const_pols, uint64_t n, uint64_t n_ext, uint64_t npols)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (!const_tree || !const_pols) return fail(GLP_E_ARG, "glp_const_tree: NULL pointer");
    if (n == 0 || npols == 0) return fail(GLP_E_ARG, "glp_const_tree: empty input");
    // everything glp_commit would reject, BEFORE the header is written: a failed call must not leave a plausible header in a
    // buffer MerkleTreeGL(Element *) would parse
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GLP_E_ARG, "glp_const_tree: sizes must be powers of two with n_ext >= n");
    if (n_ext > (1ull << 32)) return fail(GLP_E_DOMAIN, "Domain size too big for the curve");
    CU(cudaSetDevice(ctx->device));
    const uint64_t header[2] = {npols, n_ext};
    CU(cudaMemcpyAsync(const_tree, header, sizeof header, cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));                       // `header` is a stack buffer
    return glp_commit(ctx, const_tree + 2, const_tree + 2 + npols * n_ext, nullptr, const_pols, n, n_ext, npols);
}

// One FRI folding step (FRIProve::prove, src/starkpil/fri/friProve.cpp:22-108, the si > 0 branch): pol (2^pol_bits
// extension elements, three words each) -> out (2^step_bits elements).  shift_inv is the inverse coset shift of THIS
// layer (the reference squares it reduction_bits times after every step, :143-147).
extern "C" int glp_fri_fold(glp_ctx *ctx, uint64_t *out, const uint64_t *pol, uint32_t pol_bits, uint32_t step_bits, const uint64_t special_x[3],
                            uint64_t shift_inv)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (!out || !pol || !special_x) return fail(GLP_E_ARG, "glp_fri_fold: NULL pointer");
    if (pol_bits > 32) return fail(GLP_E_DOMAIN, "Domain size too big for the curve");
    if (step_bits > pol_bits || pol_bits - step_bits > FRI_MAX_RED_BITS) return fail(GLP_E_ARG, "glp_fri_fold: reduction must be between 2^0 and 2^6");
    CU(cudaSetDevice(ctx->device));
    const uint64_t n_in = 1ull << pol_bits, n_out = 1ull << step_bits;
    const gl_t *din = nullptr;
    GLP(stage_in(ctx, ctx->tmp_in, pol, (size_t)n_in * 3, &din));
    const bool out_dev = on_device(out);
    gl_t *dout = (gl_t *)out;
    if (!out_dev) { GLP(ensure(ctx, ctx->tmp_out, (size_t)n_out * 3 * sizeof(gl_t))); dout = (gl_t *)ctx->tmp_out.p; }
    FriFold a;
    a.out = dout;
    a.in = din;
    a.pol_bits = pol_bits;
    a.red_bits = pol_bits - step_bits;
    for (int d = 0; d < 3; d++) a.x[d] = special_x[d];
    a.shift_inv = shift_inv;
    a.wi = gl_canon(gl_inv(host_root(pol_bits)));
    a.wnx_inv = gl_canon(gl_inv(host_root(a.red_bits)));
    a.nx_inv = host_ninv(a.red_bits);
    const size_t smem = fri_fold_smem(a.red_bits);
    if (smem > 48 * 1024) GLP_SET_SMEM(k_fri_fold, smem);
    LAUNCH(ctx, k_fri_fold, (unsigned)((n_out + FRI_THREADS - 1) / FRI_THREADS), FRI_THREADS, smem, a);
    CU(cudaGetLastError());
    if (!out_dev) CU(cudaMemcpyAsync(out, dout, (size_t)n_out * 3 * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

// The Merkle tree of a FRI layer (friProve.cpp:110-123,252-270): the folded polynomial (pol_n = 2^pol_bits elements)
// regrouped by getTransposed into 2^next_bits rows of pol_n / 2^next_bits elements (`aux`, the tree's source), and
// MerkleTreeGL over those rows.  aux / tree / root may be NULL when not wanted.
extern "C" int glp_fri_group_tree(glp_ctx *ctx, uint64_t *aux, uint64_t *tree, uint64_t root[4], const uint64_t *pol, uint32_t pol_bits,
                                  uint32_t next_bits)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (!pol) return fail(GLP_E_ARG, "glp_fri_group_tree: pol is NULL");
    if (pol_bits > 32) return fail(GLP_E_DOMAIN, "Domain size too big for the curve");
    if (next_bits > pol_bits) return fail(GLP_E_ARG, "glp_fri_group_tree: more groups than elements");
    CU(cudaSetDevice(ctx->device));
    const uint64_t pol_n = 1ull << pol_bits, rows = 1ull << next_bits, width = 3 * (pol_n >> next_bits);
    const size_t tree_words = glp_tree_num_elements(rows);
    const gl_t *dpol = nullptr;
    GLP(stage_in(ctx, ctx->tmp_in, pol, (size_t)pol_n * 3, &dpol));
    const bool aux_dev = on_device(aux), tree_dev = on_device(tree);
    gl_t *daux = (gl_t *)aux;
    if (!aux_dev) { GLP(ensure(ctx, ctx->tmp_out, (size_t)pol_n * 3 * sizeof(gl_t))); daux = (gl_t *)ctx->tmp_out.p; }
    gl_t *dtree = (gl_t *)tree;
    if (!tree_dev) { GLP(ensure(ctx, ctx->nodes, tree_words * sizeof(gl_t))); dtree = (gl_t *)ctx->nodes.p; }
    const uint64_t nb = (pol_n + 255) / 256;
    LAUNCH_FLAT(ctx, k_fri_transpose, (unsigned)(nb > 148 * 32 ? 148 * 32 : nb), 256, daux, dpol, pol_n, next_bits);
    CU(cudaGetLastError());
    if (aux && !aux_dev) CU(cudaMemcpyAsync(aux, daux, (size_t)pol_n * 3 * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    GLP(dev_merkletree(ctx, dtree, daux, width, width, rows));
    if (tree && !tree_dev) CU(cudaMemcpyAsync(tree, dtree, tree_words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    if (root) CU(cudaMemcpyAsync(root, dtree + tree_words - 4, 4 * sizeof(gl_t), cudaMemcpyDefault, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

// ------------------------------------------------------------------------------------------------
// synthetic data, checksums, self-test
// ------------------------------------------------------------------------------------------------
extern "C" int glp_synth_fill(glp_ctx *ctx, uint64_t *dst, uint64_t nrows, uint64_t ncols, uint64_t row0, uint64_t col0,
                              uint64_t ncols_total, uint64_t seed)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (!nrows || !ncols) return GLP_OK;
    if (!dst) return fail(GLP_E_ARG, "glp_synth_fill: dst is NULL");
    CU(cudaSetDevice(ctx->device));
    const size_t words = (size_t)nrows * ncols;
    const bool dev = on_device(dst);
    gl_t *d = (gl_t *)dst;
    if (!dev) { GLP(ensure(ctx, ctx->tmp_out, words * sizeof(gl_t))); d = (gl_t *)ctx->tmp_out.p; }
    const uint64_t nb = (words + 255) / 256;
    LAUNCH_FLAT(ctx, k_synth_fill, (unsigned)(nb > 148 * 32 ? 148 * 32 : nb), 256, d, nrows, ncols, row0, col0, ncols_total, seed);
    CU(cudaGetLastError());
    if (!dev) { CU(cudaMemcpyAsync(dst, d, words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream)); CU(cudaStreamSynchronize(ctx->stream)); }
    return GLP_OK;
}

extern "C" int glp_fib_fill(glp_ctx *ctx, uint64_t *dst, uint64_t nrows, uint64_t ncols)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (!nrows || !ncols) return GLP_OK;
    if (!dst) return fail(GLP_E_ARG, "glp_fib_fill: dst is NULL");
    CU(cudaSetDevice(ctx->device));
    const size_t words = (size_t)nrows * ncols;
    const bool dev = on_device(dst);
    gl_t *d = (gl_t *)dst;
    if (!dev) { GLP(ensure(ctx, ctx->tmp_out, words * sizeof(gl_t))); d = (gl_t *)ctx->tmp_out.p; }
    const uint64_t nt = ((nrows + GLP_FIB_CHUNK - 1) / GLP_FIB_CHUNK) * ncols;
    LAUNCH_FLAT(ctx, k_fib_fill, (unsigned)((nt + 127) / 128), 128, d, nrows, ncols);
    CU(cudaGetLastError());
    if (!dev) CU(cudaMemcpyAsync(dst, d, words * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

extern "C" int glp_checksum(glp_ctx *ctx, const uint64_t *src, uint64_t n, uint64_t out[2])
{
    if (!ctx || !out) return fail(GLP_E_ARG, "glp_checksum: NULL argument");
    out[0] = out[1] = 0;
    if (!n) return GLP_OK;
    if (!on_device(src)) return fail(GLP_E_ARG, "glp_checksum: device pointer required");
    CU(cudaSetDevice(ctx->device));
    const unsigned nb = 148 * 8;
    GLP(ensure(ctx, ctx->partial, 2 * nb * sizeof(gl_t)));
    LAUNCH(ctx, k_checksum, nb, GLP_CK_THREADS, 0, (const gl_t *)src, n, (gl_t *)ctx->partial.p);
    CU(cudaGetLastError());
    std::vector<gl_t> h(2 * nb);
    CU(cudaMemcpyAsync(h.data(), ctx->partial.p, 2 * nb * sizeof(gl_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    gl_t a = 0, b = 0;
    for (unsigned i = 0; i < nb; i++) { a = gl_add(a, h[2 * i]); b = gl_add(b, h[2 * i + 1]); }
    out[0] = gl_canon(a);
    out[1] = gl_canon(b);
    return GLP_OK;
}

extern "C" int glp_field_selftest(glp_ctx *ctx, const uint64_t *a, const uint64_t *b, uint64_t n, uint64_t *out)
{
    if (!ctx || !a || !b || !out) return fail(GLP_E_ARG, "glp_field_selftest: NULL argument");
    if (!n) return GLP_OK;
    CU(cudaSetDevice(ctx->device));
    DevBuf da, db, dout;
    int rc = ensure(ctx, da, n * 8);
    if (rc == GLP_OK) rc = ensure(ctx, db, n * 8);
    if (rc == GLP_OK) rc = ensure(ctx, dout, GLP_SELFTEST_OUTS * n * 8);
    if (rc == GLP_OK) {
        cudaMemcpyAsync(da.p, a, n * 8, cudaMemcpyDefault, ctx->stream);
        cudaMemcpyAsync(db.p, b, n * 8, cudaMemcpyDefault, ctx->stream);
        LAUNCH_FLAT(ctx, k_field_selftest, (unsigned)((n + 127) / 128), 128, (const gl_t *)da.p, (const gl_t *)db.p, n, (gl_t *)dout.p);
        cudaMemcpyAsync(out, dout.p, GLP_SELFTEST_OUTS * n * 8, cudaMemcpyDefault, ctx->stream);
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = fail(GLP_E_CUDA, cudaGetErrorString(e));
    }
    release(da); release(db); release(dout);
    return rc;
}

// ------------------------------------------------------------------------------------------------
// measurement hooks
// ------------------------------------------------------------------------------------------------
extern "C" int glp_set_profiling(glp_ctx *ctx, int on)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    ctx->profiling = on != 0;
    return GLP_OK;
}
extern "C" int glp_get_profile(glp_ctx *ctx, float ms[4], uint32_t count[4])
{
    if (!ctx || !ms || !count) return fail(GLP_E_ARG, "glp_get_profile: NULL argument");
    for (int i = 0; i < 4; i++) { ms[i] = 0; count[i] = 0; }
    CU(cudaStreamSynchronize(ctx->stream));
    for (auto &sp : ctx->spans) {
        float t = 0;
        if (cudaEventElapsedTime(&t, sp.a, sp.b) == cudaSuccess && sp.kind >= 0 && sp.kind < 4) { ms[sp.kind] += t; count[sp.kind]++; }
        cudaEventDestroy(sp.a);
        cudaEventDestroy(sp.b);
    }
    ctx->spans.clear();
    cudaGetLastError();
    return GLP_OK;
}
static int measure_pipe_modes(glp_ctx *ctx, double *ops_per_s, int nmodes)
{
    CU(cudaSetDevice(ctx->device));
    GLP(ensure(ctx, ctx->partial, 2 * 148 * 8 * sizeof(gl_t)));
    cudaEvent_t a, b;
    CU(cudaEventCreate(&a));
    CU(cudaEventCreate(&b));
    const unsigned grid = 148 * 8, block = 256;
    const uint32_t iters = 4096;
    auto go = [&](int mode, uint32_t it) {
        uint64_t *sink = (uint64_t *)ctx->partial.p;
        switch (mode) {
        case 0: LAUNCH_FLAT(ctx, k_int_peak<0>, grid, block, sink, it); break;
        case 1: LAUNCH_FLAT(ctx, k_int_peak<1>, grid, block, sink, it); break;
        case 2: LAUNCH_FLAT(ctx, k_int_peak<2>, grid, block, sink, it); break;
        case 3: LAUNCH_FLAT(ctx, k_int_peak<3>, grid, block, sink, it); break;
        default: LAUNCH_FLAT(ctx, k_int_peak<4>, grid, block, sink, it); break;
        }
    };
    for (int mode = 0; mode < nmodes; mode++) {
        go(mode, 64u);                                               // warm-up
        CU(cudaEventRecord(a, ctx->stream));
        go(mode, iters);
        CU(cudaEventRecord(b, ctx->stream));
        CU(cudaEventSynchronize(b));
        float t = 0;
        CU(cudaEventElapsedTime(&t, a, b));
        const double per_thread = (double)iters * 8 * 8 * (mode == 2 ? 2 : 1);
        ops_per_s[mode] = per_thread * grid * block / (t * 1e-3);
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    return GLP_OK;
}
extern "C" int glp_measure_int_peak(glp_ctx *ctx, double ops_per_s[3])
{
    if (!ctx || !ops_per_s) return fail(GLP_E_ARG, "glp_measure_int_peak: NULL argument");
    return measure_pipe_modes(ctx, ops_per_s, 3);
}
extern "C" int glp_measure_pipe_peaks(glp_ctx *ctx, double ops_per_s[5])
{
    if (!ctx || !ops_per_s) return fail(GLP_E_ARG, "glp_measure_pipe_peaks: NULL argument");
    return measure_pipe_modes(ctx, ops_per_s, 5);
}

// ------------------------------------------------------------------------------------------------
// batched group proofs without any allocation (glp_merkle_group_proofs_static)
// ------------------------------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  glp_merkle_group_proofs stages the query
// indices in ctx->tmp_out, which ensure() may cudaMalloc / cudaFree; here that scratch is a file-scope device array of a
// fixed size, GLP_GROUP_PROOFS_STATIC_MAX indices (32 KiB), which every GPU that loads the library holds until exit.
// In host code `g_gp_static_idx` is only the host-side stand-in for the symbol, never a device address: the address the
// kernel gets comes from cudaGetSymbolAddress, asked once per device with that device current (gp_static_scratch).
// Every call on a device shares the array and the array keeps the last call's indices; a call overwrites the first nidx
// entries and the kernel reads only those, so nothing has to start at zero -- but two calls that overlap on the GPU
// (different streams, different contexts of one device) race on it.
static __device__ __align__(256) unsigned long long g_gp_static_idx[GLP_GROUP_PROOFS_STATIC_MAX];

namespace {
constexpr int GP_STATIC_MAX_DEVICES = 64;
std::mutex g_gp_static_mu;
uint64_t *g_gp_static_addr[GP_STATIC_MAX_DEVICES];     // per device: the device address of g_gp_static_idx

int gp_static_scratch(int device, uint64_t **out)      // the caller has made `device` current
{
    if (device < 0 || device >= GP_STATIC_MAX_DEVICES) return fail(GLP_E_ARG, "glp_merkle_group_proofs_static: device index out of range");
    std::lock_guard<std::mutex> lk(g_gp_static_mu);
    if (!g_gp_static_addr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_gp_static_idx);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(GLP_E_CUDA, std::string("glp_merkle_group_proofs_static: cudaGetSymbolAddress failed (") + cudaGetErrorString(e) + ")");
        }
        g_gp_static_addr[device] = (uint64_t *)p;
    }
    *out = g_gp_static_addr[device];
    return GLP_OK;
}
}  // namespace

extern "C" int glp_merkle_group_proofs_static(glp_ctx *ctx, uint64_t *proofs, const uint64_t *source, const uint64_t *nodes, uint64_t width,
                                              uint64_t height, const uint64_t *idx, uint64_t nidx)
{
    if (!ctx) return fail(GLP_E_ARG, "ctx is NULL");
    if (nidx > GLP_GROUP_PROOFS_STATIC_MAX)
        return fail(GLP_E_ARG, "glp_merkle_group_proofs_static: " + std::to_string(nidx) + " indices exceed the static scratch (" +
                                   std::to_string((unsigned long long)GLP_GROUP_PROOFS_STATIC_MAX) + "); use glp_merkle_group_proofs");
    CU(cudaSetDevice(ctx->device));
    uint64_t *didx = nullptr;
    GLP(gp_static_scratch(ctx->device, &didx));
    if (nidx == 0) return GLP_OK;                                // set-up only: lets a caller do it before a capture begins
    if (!proofs || !nodes || !idx || (!source && width)) return fail(GLP_E_ARG, "glp_merkle_group_proofs_static: NULL pointer");
    if (!is_pow2(height)) return fail(GLP_E_ARG, "glp_merkle_group_proofs_static: height must be a power of two");
    if (on_device(idx)) return fail(GLP_E_ARG, "glp_merkle_group_proofs_static: idx is a host array");
    if (!on_device(proofs) || !on_device(nodes) || (width && !on_device(source)))
        return fail(GLP_E_ARG, "glp_merkle_group_proofs_static: proofs, source and nodes must be device memory (no staging buffer is allocated)");
    for (uint64_t q = 0; q < nidx; q++)
        if (idx[q] >= height) return fail(GLP_E_ARG, "glp_merkle_group_proofs_static: index out of range");
    const uint64_t words = width + 4 * ilog2_u64(height);
    CU(cudaMemcpyAsync(didx, idx, nidx * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
    LAUNCH(ctx, k_group_proofs, (unsigned)nidx, 128, 0, (gl_t *)proofs, words, (const gl_t *)source, (const gl_t *)nodes, width, height,
           (const uint64_t *)didx, nidx);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(ctx->stream));
    return GLP_OK;
}

// ------------------------------------------------------------------------------------------------
// several devices, one process (glp_multi_*)
// ------------------------------------------------------------------------------------------------
#include "multi.cuh"
