// Stand-alone stand-in for the one type the drop-in headers need from the reference's
// goldilocks_base_field.hpp (src/goldilocks/src/goldilocks_base_field.hpp:20-27): Goldilocks::Element, a
// struct holding one uint64_t, plus the handful of conversions the wrapper tests use.
//
// Inside the reference tree this file is never used: ntt_goldilocks.hpp / merkleTreeGL.hpp include the real
// goldilocks_base_field.hpp there (its include guard GOLDILOCKS_BASE wins).  It exists so that the headers,
// and tests/cpp/test_dropin.cpp, compile without the reference on the GPU box.
#ifndef GOLDILOCKS_BASE
#define GOLDILOCKS_BASE
#include <stdint.h>

#define GOLDILOCKS_PRIME 0xFFFFFFFF00000001ULL

class Goldilocks
{
public:
    typedef struct
    {
        uint64_t fe;
    } Element;

    static inline Element fromU64(uint64_t v)
    {
        Element e;
        e.fe = v >= GOLDILOCKS_PRIME ? v - GOLDILOCKS_PRIME : v;
        return e;
    }
    static inline uint64_t toU64(const Element &e) { return e.fe >= GOLDILOCKS_PRIME ? e.fe - GOLDILOCKS_PRIME : e.fe; }
    static inline Element zero() { return fromU64(0); }
    static inline Element one() { return fromU64(1); }
};
#endif
