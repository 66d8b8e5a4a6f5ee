/* TEST INFRASTRUCTURE ONLY (oracle build shim) -- not part of the product.
 *
 * Minimal stand-in for <gmp.h> so that the reference's own sources
 * (/root/reference/src/goldilocks/src/ntt_goldilocks.hpp:61-135) compile in an
 * image that ships libgmp.so.10 but not the GMP development headers.
 * Every value the reference ever puts into an mpz_t on this path is < 2^128
 * (the modulus p, (p-1)/2, a non-residue, the inverse of 2), so the dozen
 * mpz_* calls are implemented on unsigned __int128 with no library at all.
 */
#ifndef GLP_ORACLE_GMP_SHIM_H
#define GLP_ORACLE_GMP_SHIM_H
#include <stddef.h>
#include <stdint.h>

typedef unsigned __int128 glp_u128;
typedef struct { glp_u128 v; } __mpz_struct;
typedef __mpz_struct mpz_t[1];

static inline void mpz_init(mpz_t x) { x->v = 0; }
static inline void mpz_clear(mpz_t x) { (void)x; }
static inline void mpz_set(mpz_t r, const mpz_t a) { r->v = a->v; }
static inline void mpz_set_ui(mpz_t r, unsigned long a) { r->v = a; }
static inline unsigned long mpz_get_ui(const mpz_t a) { return (unsigned long)a->v; }
static inline void mpz_add_ui(mpz_t r, const mpz_t a, unsigned long b) { r->v = a->v + b; }
static inline void mpz_fdiv_q_2exp(mpz_t r, const mpz_t a, unsigned long b) { r->v = a->v >> b; }
static inline int mpz_tstbit(const mpz_t a, unsigned long bit) { return (int)((a->v >> bit) & 1); }
static inline int mpz_cmp_ui(const mpz_t a, unsigned long b) { return a->v < b ? -1 : (a->v > b ? 1 : 0); }
/* only the (count=1, order=1, size=8, endian=0, nails=0) form is used by the reference */
static inline void mpz_import(mpz_t r, size_t count, int order, size_t size, int endian, size_t nails, const void *op)
{
    (void)count; (void)order; (void)size; (void)endian; (void)nails;
    r->v = *(const uint64_t *)op;
}
static inline glp_u128 glp_mulmod_u128(glp_u128 a, glp_u128 b, glp_u128 m)
{
    /* operands < 2^64 on this path, so the product fits 128 bits */
    return (a * b) % m;
}
static inline void mpz_powm(mpz_t r, const mpz_t base, const mpz_t exp, const mpz_t mod)
{
    glp_u128 b = base->v % mod->v, e = exp->v, acc = 1;
    while (e) {
        if (e & 1) acc = glp_mulmod_u128(acc, b, mod->v);
        b = glp_mulmod_u128(b, b, mod->v);
        e >>= 1;
    }
    r->v = acc;
}
/* modulus is the prime p on this path: a^-1 = a^(p-2) */
static inline int mpz_invert(mpz_t r, const mpz_t a, const mpz_t mod)
{
    mpz_t e; e->v = mod->v - 2;
    mpz_powm(r, a, e, mod);
    return 1;
}
#endif
