// Test infrastructure (tests/test_oracle_kat.py compiles and runs it): the division ops.cuh SharedDiv performs for
// Ramp.js:37 (t / duration, many t by one duration) against the hardware's IEEE division.  Markstein's step: with
// y = RN(1/b), q = RN(a*y), r = a - b*q (one FMA, exact), RN(q + r*y) == RN(a/b).
// usage: kat_div [millions of cases per class = 150]
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
static uint64_t s = 0x9E3779B97F4A7C15ull;
static uint64_t rnd(void) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }
static double from_bits(uint64_t u) { double d; memcpy(&d, &u, 8); return d; }
static int check(double a, double b) {
  const double y = 1.0 / b, q = a * y, r = fma(-q, b, a), q2 = fma(r, y, q), w = a / b;
  return memcmp(&q2, &w, 8) == 0;
}
int main(int argc, char **argv) {
  long long bad = 0, n = 0;
  const long long M = (argc > 1 ? atoll(argv[1]) : 150) * 1000000LL;
  // (1) integers as Ramp sees them: t in [0, 2^32), duration in [1, 2^32)
  for (long long i = 0; i < M; i++) {
    double a = (double)(rnd() >> (32 + (rnd() & 15))), b = (double)((rnd() >> (32 + (rnd() & 15))) | 1);
    n++; if (!check(a, b)) { if (bad < 5) printf("int fail %.17g / %.17g\n", a, b); bad++; }
  }
  // (2) arbitrary doubles with exponents in [-200, 200], random significands, both signs
  for (long long i = 0; i < M; i++) {
    uint64_t ua = (rnd() & 0x800FFFFFFFFFFFFFull) | ((uint64_t)(1023 - 200 + (rnd() % 401)) << 52);
    uint64_t ub = (rnd() & 0x800FFFFFFFFFFFFFull) | ((uint64_t)(1023 - 200 + (rnd() % 401)) << 52);
    n++; if (!check(from_bits(ua), from_bits(ub))) { if (bad < 5) printf("fp fail %a / %a\n", from_bits(ua), from_bits(ub)); bad++; }
  }
  // (3) hard significands: all ones, one below/above a power of two, against random and hard numerators
  const uint64_t hard[] = {0x000FFFFFFFFFFFFFull, 0x000FFFFFFFFFFFFEull, 0x0000000000000001ull, 0x0000000000000000ull, 0x0008000000000000ull, 0x0007FFFFFFFFFFFFull};
  for (int hb = 0; hb < 6; hb++)
    for (long long i = 0; i < M / 8 + 1000; i++) {
      uint64_t ub = hard[hb] | (1023ull << 52);
      uint64_t ua = (i & 7) < 6 ? (hard[i & 7 ? (i & 7) - 1 : 0] | (1023ull << 52)) ^ (rnd() & 0xFF) : (rnd() & 0x000FFFFFFFFFFFFFull) | (1023ull << 52);
      n++; if (!check(from_bits(ua), from_bits(ub))) { if (bad < 5) printf("hard fail %a / %a\n", from_bits(ua), from_bits(ub)); bad++; }
    }
  printf("%lld cases, %lld mismatches\n", n, bad);
  return bad != 0;
}
