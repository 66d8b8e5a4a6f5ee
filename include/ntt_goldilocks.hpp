// Drop-in replacement for the reference's src/goldilocks/src/ntt_goldilocks.hpp (class NTT_Goldilocks,
// :12-174) and ntt_goldilocks.cpp.  Same class name, constructor and public method signatures, so
// src/starkpil/starks.hpp:46-47 (`NTT_Goldilocks ntt; NTT_Goldilocks nttExtended;`), the Stark commit stages
// (src/starkpil/starks.cpp:53-57,134-138,215-219,262-291) and FRI (src/starkpil/fri/friProve.cpp:93-105)
// compile unchanged -- but every transform runs on the B200 through the C ABI in glprover.h.
// Header-only: link with -lglprover.  No CPU implementation: if the CUDA library cannot
// run, the methods throw std::runtime_error.
//
// Differences a caller can observe (all documented in DESIGN.md "Boundary"):
//  * outputs are canonical (< p); the reference may leave non-canonical representatives.  Equal under
//    Goldilocks::toU64, which is what every reference test compares (tests/tests.cpp).
//  * buffer / nphase / nblock are accepted and ignored (performance hints in the reference; results are
//    identical for every value, tests.cpp:2316-2338).
//  * pointers may be host or device memory.
#ifndef NTT_GOLDILOCKS
#define NTT_GOLDILOCKS

#if defined(__has_include)
#if __has_include("goldilocks_base_field.hpp")
#include "goldilocks_base_field.hpp"
#else
#include "glp_goldilocks_element.hpp"
#endif
#else
#include "goldilocks_base_field.hpp"
#endif
#include "glp_cxx.hpp"
#include <sys/types.h>
#include <vector>

#define NUM_PHASES 3
#define NUM_BLOCKS 1

class NTT_Goldilocks
{
private:
    u_int32_t s = 0;        // log2 of the largest domain this object was built for (ntt_goldilocks.hpp:14,90-99)
    u_int32_t nThreads = 0; // kept for source compatibility; meaningless on the GPU
    int extension = 1;      // reversePermutation zero-fills rows >= size/extension (ntt_goldilocks.cpp:259-277)
    // root() hands out references, so the values need a home; the table is built on the first root() call
    // only (nothing on the hot path calls root(): the device keeps its own twiddle tables)
    std::vector<Goldilocks::Element> rootsTable;

    static u_int32_t log2(u_int64_t size)
    {
        u_int32_t res = 0;
        while (size > 1)
        {
            size >>= 1;
            res++;
        }
        return res;
    }
    static uint64_t *raw(Goldilocks::Element *p) { return reinterpret_cast<uint64_t *>(p); }

public:
    NTT_Goldilocks(u_int64_t maxDomainSize, u_int32_t _nThreads = 0, int extension_ = 1)
    {
        static_assert(sizeof(Goldilocks::Element) == sizeof(uint64_t), "Goldilocks::Element must be one 64-bit word");
        if (maxDomainSize == 0)
            return; // inert object (ntt_goldilocks.hpp:55-56); no GPU work, no context
        nThreads = _nThreads;
        extension = extension_;
        const u_int32_t domainPow = log2(maxDomainSize);
        if (domainPow > 32)
            throw std::range_error("Domain size too big for the curve"); // ntt_goldilocks.hpp:94-97
        s = domainPow == 0 ? 1 : domainPow;
        // twiddle tables: built on the device by the first transform of each size and cached in the
        // thread's context (the reference rebuilds 2^s sequential products per object, :121-124)
    }
    ~NTT_Goldilocks() {}

    // shift^i and shift^i/N tables (ntt_goldilocks.hpp:153-165) live in the library's cache, keyed by size,
    // so the reference's stale-table problem (computed once for the first N) cannot occur
    inline void computeR(int /*N*/) {}

    void NTT(Goldilocks::Element *dst, Goldilocks::Element *src, u_int64_t size, u_int64_t ncols = 1, Goldilocks::Element *buffer = NULL,
             u_int64_t nphase = NUM_PHASES, u_int64_t nblock = NUM_BLOCKS, bool inverse = false, bool extend = false)
    {
        (void)buffer;
        (void)nphase;
        (void)nblock;
        if (ncols == 0 || size == 0)
            return; // ntt_goldilocks.cpp:168-171
        // an object built with extension > 1 reads rows >= size/extension as zero (reversePermutation's zero-fill,
        // ntt_goldilocks.cpp:259-277); in place that combination asserts in the reference (:301-304)
        const uint64_t live = extension > 1 ? size / (uint64_t)extension : size;
        if (extension > 1 && (dst == src || dst == NULL))
            throw std::invalid_argument("NTT_Goldilocks::NTT: in place with extension > 1 is not implemented (ntt_goldilocks.cpp:301-304)");
        glp::check(glp_ntt_live(glp::thread_ctx(), raw(dst), raw(src), size, ncols, inverse ? 1 : 0, extend ? 1 : 0, live), "NTT_Goldilocks::NTT");
    }

    void INTT(Goldilocks::Element *dst, Goldilocks::Element *src, u_int64_t size, u_int64_t ncols = 1, Goldilocks::Element *buffer = NULL,
              u_int64_t nphase = NUM_PHASES, u_int64_t nblock = NUM_BLOCKS, bool extend = false)
    {
        if (ncols == 0 || size == 0)
            return;
        NTT(dst == NULL ? src : dst, src, size, ncols, buffer, nphase, nblock, true, extend); // ntt_goldilocks.cpp:315-322
    }

    void reversePermutation(Goldilocks::Element *dst, Goldilocks::Element *src, u_int64_t size, u_int64_t offset_cols, u_int64_t ncols,
                            u_int64_t ncols_all)
    {
        glp::check(glp_reverse_permutation(glp::thread_ctx(), raw(dst), raw(src), size, offset_cols, ncols, ncols_all, (uint64_t)extension),
                   "NTT_Goldilocks::reversePermutation");
    }

    // W[domainPow]^idx (ntt_goldilocks.hpp:169-172)
    inline Goldilocks::Element &root(u_int32_t domainPow, u_int64_t idx)
    {
        if (rootsTable.empty())
        {
            rootsTable.resize((size_t)1 << s);
            const unsigned __int128 p = 0xFFFFFFFF00000001ULL;
            const uint64_t w = glp_root(s, 1);
            uint64_t acc = 1;
            for (size_t i = 0; i < rootsTable.size(); i++)
            {
                rootsTable[i].fe = acc;
                acc = (uint64_t)(((unsigned __int128)acc * w) % p);
            }
        }
        return rootsTable[idx << (s - domainPow)];
    }

    // output[i,c] = P_c(shift * w_ext^i) (ntt_goldilocks.cpp:326-352); output may alias input
    void extendPol(Goldilocks::Element *output, Goldilocks::Element *input, uint64_t N_Extended, uint64_t N, uint64_t ncols,
                   Goldilocks::Element *buffer = NULL, u_int64_t nphase = NUM_PHASES, u_int64_t nblock = NUM_BLOCKS)
    {
        (void)buffer;
        (void)nphase;
        (void)nblock;
        if (glp::multi_enabled())
            glp::check(glp_multi_extend_pol(glp::thread_multi(), raw(output), raw(input), N_Extended, N, ncols), "NTT_Goldilocks::extendPol");
        else
            glp::check(glp_extend_pol(glp::thread_ctx(), raw(output), raw(input), N_Extended, N, ncols), "NTT_Goldilocks::extendPol");
    }
};

#endif
