// What the UNCHANGED folding loop of FRIProve::prove costs through the header swap (src/starkpil/fri/friProve.cpp:93-105: an
// NTT_Goldilocks(nX, 1) per group and ntt.INTT(ppar_c, ppar, nX, 3) on nX = 32 points), per call, next to one glp_fri_fold of the
// same step.  g++ -std=c++17 -O2 -I../../include fri_unchanged_probe.cpp -L../../xlayer_prover_b200 -lglprover -Wl,-rpath,...
#include "ntt_goldilocks.hpp"
#include <chrono>
#include <cstdio>
#include <vector>
int main()
{
    const uint64_t nX = 32, groups = 1 << 14, polBits = 19;      // timed on 2^14 of the 2^19 groups of the first zkEVM step
    std::vector<Goldilocks::Element> pol((size_t)3 << polBits), ppar(nX * 3), ppar_c(nX * 3);
    for (size_t i = 0; i < pol.size(); i++) pol[i].fe = i * 0x9E3779B97F4A7C15ULL % 0xFFFFFFFF00000001ULL;
    { NTT_Goldilocks warm(nX, 1); warm.INTT(ppar_c.data(), ppar.data(), nX, 3); }
    auto t0 = std::chrono::steady_clock::now();
    for (uint64_t g = 0; g < groups; g++) {
        for (uint64_t i = 0; i < nX; i++) for (int k = 0; k < 3; k++) ppar[i * 3 + k] = pol[((i * groups + g) * 3 + k) % pol.size()];
        NTT_Goldilocks ntt(nX, 1);
        ntt.INTT(ppar_c.data(), ppar.data(), nX, 3);
    }
    const double per_call = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / groups;
    std::vector<uint64_t> out((size_t)3 << (polBits - 5));
    const uint64_t x[3] = {5, 6, 7};
    glp_fri_fold(glp::thread_ctx(), out.data(), (const uint64_t *)pol.data(), polBits, polBits - 5, x, 123);
    t0 = std::chrono::steady_clock::now();
    glp_fri_fold(glp::thread_ctx(), out.data(), (const uint64_t *)pol.data(), polBits, polBits - 5, x, 123);
    const double fold = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("unchanged loop: %.1f us per INTT(32 x 3) call through the drop-in -> %.2f s for the 2^19 groups of one 2^24 -> 2^19 step (one host thread);\n"
           "glp_fri_fold of a 2^19 -> 2^14 step with host buffers: %.2f ms\n", per_call * 1e6, per_call * (1 << 19), fold * 1e3);
    return 0;
}
