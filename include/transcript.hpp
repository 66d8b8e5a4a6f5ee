// Drop-in replacement for the reference's src/starkpil/transcript/transcript.{hpp,cpp} (class Transcript): the
// Fiat-Shamir sponge of the STARK.  Same public members and methods; the one arithmetic step -- the width-12
// Poseidon permutation PoseidonGoldilocks::hash_full_result (transcript.cpp:24,50) -- runs on the B200 through
// glp_hash_full_result, so a prover that links libglprover.so needs no CPU Poseidon at all.  The cursor logic
// below restates transcript.cpp:4-91 line by line (put/_add1 :4-30, getField :32-38, getFields1 :40-59,
// getPermutations :61-91).  Header-only: link with -lglprover.
#ifndef TRANSCRIPT_CLASS
#define TRANSCRIPT_CLASS

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
#include <cstring>
#include <vector>

#define TRANSCRIPT_STATE_SIZE 4
#define TRANSCRIPT_PENDING_SIZE 8
#define TRANSCRIPT_OUT_SIZE 12

class Transcript
{
private:
    void permute()
    {
        Goldilocks::Element inputs[TRANSCRIPT_OUT_SIZE];
        std::memcpy(inputs, pending, TRANSCRIPT_PENDING_SIZE * sizeof(Goldilocks::Element));
        std::memcpy(&inputs[TRANSCRIPT_PENDING_SIZE], state, TRANSCRIPT_STATE_SIZE * sizeof(Goldilocks::Element));
        glp::check(glp_hash_full_result(glp::thread_ctx(), reinterpret_cast<uint64_t *>(out), reinterpret_cast<const uint64_t *>(inputs), 1),
                   "Transcript: hash_full_result");
        out_cursor = TRANSCRIPT_OUT_SIZE;
        std::memset(pending, 0, TRANSCRIPT_PENDING_SIZE * sizeof(Goldilocks::Element));
        pending_cursor = 0;
        std::memcpy(state, out, TRANSCRIPT_STATE_SIZE * sizeof(Goldilocks::Element));
    }
    void _add1(Goldilocks::Element input)
    {
        pending[pending_cursor] = input;
        pending_cursor++;
        out_cursor = 0;
        if (pending_cursor == TRANSCRIPT_PENDING_SIZE)
            permute();
    }

public:
    Goldilocks::Element state[TRANSCRIPT_STATE_SIZE];
    Goldilocks::Element pending[TRANSCRIPT_PENDING_SIZE];
    Goldilocks::Element out[TRANSCRIPT_OUT_SIZE];

    unsigned int pending_cursor = 0;
    unsigned int out_cursor = 0;
    unsigned int state_cursor = 0;

    Transcript()
    {
        std::memset(state, 0, sizeof state);
        std::memset(pending, 0, sizeof pending);
        std::memset(out, 0, sizeof out);
    }
    void put(Goldilocks::Element *input, uint64_t size)
    {
        for (uint64_t i = 0; i < size; i++)
            _add1(input[i]);
    }
    Goldilocks::Element getFields1()
    {
        if (out_cursor == 0)
            permute();
        Goldilocks::Element res = out[(TRANSCRIPT_OUT_SIZE - out_cursor) % TRANSCRIPT_OUT_SIZE];
        out_cursor--;
        return res;
    }
    void getField(Goldilocks::Element *output)
    {
        for (int i = 0; i < 3; i++)
            output[i] = getFields1();
    }
    void getPermutations(uint64_t *res, uint64_t n, uint64_t nBits)
    {
        const uint64_t totalBits = n * nBits;
        const uint64_t NFields = (totalBits - 1) / 63 + 1;
        std::vector<Goldilocks::Element> fields(NFields);
        for (uint64_t i = 0; i < NFields; i++)
            fields[i] = getFields1();
        uint64_t curField = 0, curBit = 0;
        for (uint64_t i = 0; i < n; i++)
        {
            uint64_t a = 0;
            for (uint64_t j = 0; j < nBits; j++)
            {
                const uint64_t bit = (Goldilocks::toU64(fields[curField]) >> curBit) & 1;
                if (bit)
                    a = a + (1ull << j);
                curBit++;
                if (curBit == 63)
                {
                    curBit = 0;
                    curField++;
                }
            }
            res[i] = a;
        }
    }
};

#endif
