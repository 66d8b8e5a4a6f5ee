/* TEST INFRASTRUCTURE ONLY (oracle build shim) -- not part of the product.
 *
 * Minimal stand-in for <gmpxx.h>: the reference uses mpz_class only in string /
 * signed conversions (/root/reference/src/goldilocks/src/goldilocks_base_field_tools.hpp:120-237),
 * never on the hot path.  Backed by a signed __int128.  The real header also
 * drags in <cstring> and <iostream>, which the reference relies on.
 */
#ifndef GLP_ORACLE_GMPXX_SHIM_H
#define GLP_ORACLE_GMPXX_SHIM_H
#include <cstdint>
#include <cstring>
#include <iostream>
#include <string>
#include <algorithm>
#include "gmp.h"

class mpz_class
{
public:
    __int128 v;
    mpz_class() : v(0) {}
    mpz_class(const mpz_class &o) = default;
    mpz_class(uint64_t a) : v((__int128)a) {}
    mpz_class(int a) : v(a) {}
    mpz_class(long a) : v(a) {}
    mpz_class(long long a) : v(a) {}
    mpz_class(const std::string &s, int radix = 10) : v(0)
    {
        size_t i = 0; bool neg = false;
        if (i < s.size() && (s[i] == '-' || s[i] == '+')) { neg = (s[i] == '-'); i++; }
        if (radix == 16 && i + 1 < s.size() && s[i] == '0' && (s[i + 1] == 'x' || s[i + 1] == 'X')) i += 2;
        for (; i < s.size(); i++) {
            char c = s[i]; int d;
            if (c >= '0' && c <= '9') d = c - '0';
            else if (c >= 'a' && c <= 'z') d = c - 'a' + 10;
            else if (c >= 'A' && c <= 'Z') d = c - 'A' + 10;
            else break;
            v = v * radix + d;
        }
        if (neg) v = -v;
    }
    mpz_class &operator=(const mpz_class &o) = default;
    uint64_t get_ui() const { return (uint64_t)(v < 0 ? -v : v); }
    int64_t get_si() const { return (int64_t)v; }
    std::string get_str(int radix = 10) const
    {
        if (v == 0) return "0";
        unsigned __int128 a = v < 0 ? (unsigned __int128)(-v) : (unsigned __int128)v;
        std::string out;
        while (a) { int d = (int)(a % radix); out.push_back(d < 10 ? '0' + d : 'a' + d - 10); a /= radix; }
        if (v < 0) out.push_back('-');
        std::reverse(out.begin(), out.end());
        return out;
    }
};
inline mpz_class operator+(const mpz_class &a, const mpz_class &b) { mpz_class r; r.v = a.v + b.v; return r; }
inline mpz_class operator-(const mpz_class &a, const mpz_class &b) { mpz_class r; r.v = a.v - b.v; return r; }
/* GMP's % truncates toward zero like C; the reference always adds p first */
inline mpz_class operator%(const mpz_class &a, const mpz_class &b) { mpz_class r; r.v = a.v % b.v; return r; }
inline bool operator>(const mpz_class &a, const mpz_class &b) { return a.v > b.v; }
inline bool operator<(const mpz_class &a, const mpz_class &b) { return a.v < b.v; }
inline bool operator==(const mpz_class &a, const mpz_class &b) { return a.v == b.v; }
#endif
