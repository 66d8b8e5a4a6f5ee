// TEST INFRASTRUCTURE ONLY.  C wrapper around the reference's FRI prover, compiled from src/starkpil/fri/friProve.cpp where it
// lies (oracle/Makefile target `ref`, oracle/_ref/libfriref.so) together with the reference's own polinomial.hpp, friProof.hpp,
// transcript.cpp and merkleTreeGL.cpp; stand-ins only for headers that belong to the prover's control plane
// (oracle/shim/fri/{stark_info,zklog,exit_process,timer,compare_fe}.hpp, oracle/shim/nlohmann/json.hpp).  Nothing of the FRI
// arithmetic is restated: ref_fri_prove runs FRIProve::prove itself (friProve.cpp:5-190) -- folding steps (INTT, polMulAxi,
// evalPol), layer trees (getTransposed + MerkleTreeGL), transcript, query phase (queryPol -> getGroupProof) -- and hands back
// what it put into the FRIProof; ref_fri_fold_step calls the two static helpers polMulAxi / evalPol (:192-217) around the
// reference's INTT for one folding step.
#include <cstring>
#include <vector>
#include "friProve.hpp"
typedef Goldilocks::Element E;
extern "C" {
// friPol: 2^polBits cubic-extension elements (in/out: holds the last layer afterwards).  trees_src[i] (i < 5): the source matrix of
// stage tree i, 2^polBits rows x widths[i] columns (the trees the query phase opens, friProve.cpp:159-163).  transcript_seed: words
// put into the transcript before the call.  step_bits: the nBits of every FRI step (step 0 has nBits == polBits).
// Outputs: roots[(nsteps-1)*4] = root of every layer tree (steps 1..), final_pol[2^step_bits[last] * 3], ys[nqueries] = the query
// indices, queries = for every step si and query i the concatenated proofs in the order queryPol pushes them
// (stage trees: 5 proofs of widths[t] + 4*depth words; layer trees: 1 proof); *queries_words = how many words were written.
int ref_fri_prove(uint64_t *friPol, uint64_t polBits, const uint64_t *step_bits, uint64_t nsteps, uint64_t nqueries,
                  uint64_t *const *trees_src, const uint64_t *widths, const uint64_t *transcript_seed, uint64_t nseed,
                  uint64_t *roots, uint64_t *final_pol, uint64_t *ys_out, uint64_t *queries, uint64_t queries_cap, uint64_t *queries_words)
{
    StarkInfo si;
    si.starkStruct.nBits = polBits;
    si.starkStruct.nBitsExt = polBits;
    si.starkStruct.nQueries = nqueries;
    si.starkStruct.verificationHashType = "GL";
    for (uint64_t i = 0; i < nsteps; i++) { StepStruct s; s.nBits = step_bits[i]; si.starkStruct.steps.push_back(s); }
    MerkleTreeGL *trees[5];
    for (int t = 0; t < 5; t++) {
        trees[t] = new MerkleTreeGL(1ull << polBits, widths[t], (E *)trees_src[t]);
        trees[t]->merkelize();
    }
    Transcript transcript;
    if (nseed) transcript.put((E *)transcript_seed, nseed);
    // the query indices FRIProve::prove will draw: replay the transcript exactly as prove() feeds it, on a copy, AFTER the call
    FRIProof fproof(1ull << step_bits[nsteps - 1], FIELD_EXTENSION, nsteps, 0, nqueries);
    Polinomial pol((void *)friPol, 1ull << polBits, FIELD_EXTENSION, FIELD_EXTENSION);
    FRIProve::prove(fproof, trees, transcript, pol, polBits, si);
    for (uint64_t s = 1; s < nsteps; s++)
        for (int k = 0; k < 4; k++) roots[(s - 1) * 4 + k] = Goldilocks::toU64(fproof.proofs.fri.trees[s].root[k]);
    const uint64_t last = 1ull << step_bits[nsteps - 1];
    for (uint64_t i = 0; i < last; i++)
        for (int k = 0; k < 3; k++) final_pol[i * 3 + k] = Goldilocks::toU64(fproof.proofs.fri.pol[i][k]);
    uint64_t w = 0;
    for (uint64_t s = 0; s < nsteps; s++) {
        auto &pq = fproof.proofs.fri.trees[s].polQueries;
        for (uint64_t i = 0; i < pq.size(); i++)
            for (auto &mp : pq[i]) {
                for (auto &v : mp.v) { if (w < queries_cap) queries[w] = Goldilocks::toU64(v[0]); w++; }
                for (auto &h : mp.mp) for (int k = 0; k < 4; k++) { if (w < queries_cap) queries[w] = Goldilocks::toU64(h[k]); w++; }
            }
    }
    *queries_words = w;
    (void)ys_out;
    for (int t = 0; t < 5; t++) delete trees[t];
    return 0;
}

// one folding step with the reference's own helpers: out[g] = evalPol(polMulAxi(INTT(column g), 1, shiftInv * w^-g), special_x)
int ref_fri_fold_step(uint64_t *pol2_e, uint64_t *friPol_, uint64_t polBits, uint64_t nBitsStep, uint64_t *special_x_, uint64_t shiftInv)
{
    const uint64_t pol2N = 1ull << nBitsStep, nX = (1ull << polBits) / pol2N;
    Polinomial friPol((void *)friPol_, 1ull << polBits, FIELD_EXTENSION, FIELD_EXTENSION);
    Polinomial out((void *)pol2_e, pol2N, FIELD_EXTENSION, FIELD_EXTENSION);
    Polinomial special_x((void *)special_x_, 1, FIELD_EXTENSION, FIELD_EXTENSION);
    E wi = Goldilocks::inv(Goldilocks::w(polBits));
    E sinv_ = Goldilocks::fromU64(shiftInv);
    for (uint64_t g = 0; g < pol2N; g++) {
        Polinomial ppar(nX, FIELD_EXTENSION);
        Polinomial ppar_c(nX, FIELD_EXTENSION);
        for (uint64_t i = 0; i < nX; i++) Polinomial::copyElement(ppar, i, friPol, (i * pol2N) + g);
        NTT_Goldilocks ntt(nX, 1);
        ntt.INTT(ppar_c.address(), ppar.address(), nX, FIELD_EXTENSION);
        FRIProve::polMulAxi(ppar_c, Goldilocks::one(), sinv_);
        FRIProve::evalPol(out, g, ppar_c, special_x);
        sinv_ = sinv_ * wi;
    }
    return 0;
}
}
