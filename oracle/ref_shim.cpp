// TEST INFRASTRUCTURE. C wrappers around the REAL reference functions math::fac/comb/perm/pow
// (declared in /root/reference/workspace/lib/math/include/math/Combinatorics.h:12-19), compiled
// together with the reference's own Combinatorics.cpp into oracle/_ref/. No reference source is
// copied: this file only declares and forwards.
#include <cstdint>
#include <stdexcept>
namespace math {
uint64_t fac(uint64_t n);
uint64_t comb(uint64_t n, uint64_t k);
uint64_t perm(uint64_t n, uint64_t k);
template <typename T, typename U> T pow(T base, U exp);
}
extern "C" {
// returns 0 and sets *threw when the reference throws (fac(21): CombinatoricsTest.cpp)
uint64_t ref_fac(uint64_t n, int* threw) {
    *threw = 0;
    try { return math::fac(n); } catch (const std::runtime_error&) { *threw = 1; return 0; }
}
uint64_t ref_comb(uint64_t n, uint64_t k) { return math::comb(n, k); }
uint64_t ref_perm(uint64_t n, uint64_t k) { return math::perm(n, k); }
double ref_pow(double b, uint64_t e) { return math::pow<double, uint64_t>(b, e); }
}
