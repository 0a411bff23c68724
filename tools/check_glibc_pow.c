/* Validates csrc/ks_glibc_pow.h against the running libm: every pair of a positive finite x and a finite non-zero y must
 * give the bit pattern pow() gives, or be declined inside the one unported band (512 <= |y log x| < 1024).  gcc -O2 -ffp-contract=off tools/check_glibc_pow.c -lm && ./a.out [n] */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#define __device__
static inline uint64_t bits_of(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }
static inline double dbl_of(uint64_t u) { double d; memcpy(&d, &u, 8); return d; }
#define KS_PF_FMA(a, b, c) __builtin_fma((a), (b), (c))
#define KS_PF_ADD(a, b) ((a) + (b))
#define KS_PF_SUB(a, b) ((a) - (b))
#define KS_PF_MUL(a, b) ((a) * (b))
#define KS_PF_BITS(d) bits_of(d)
#define KS_PF_DBL(u) dbl_of(u)
#define KS_PF_I2D(i) ((double)(i))
#define KS_PF_FN static inline
#include "../funsearch_kubernetes_simulator_b200/csrc/ks_glibc_pow_tables.inc"
#include "../funsearch_kubernetes_simulator_b200/csrc/ks_glibc_pow.h"
#include "../funsearch_kubernetes_simulator_b200/csrc/ks_glibc_explog.h"

static uint64_t rng = 88172645463325252ull;
static uint64_t next(void) { rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17; return rng; }
static double uni(void) { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }

int main(int argc, char** argv) {
  long n = argc > 1 ? atol(argv[1]) : 20000000, bad = 0, fast = 0;
  const double ys[] = {0.5, 0.6, 2.0, 2.5, 3.0, 1.5, 0.25, -1.0, -0.5, 1e-3, 7.0, 0.3333333333333333};
  for (long it = 0; it < n; ++it) {
    double x, y;
    switch (it & 7) {
      case 0: x = uni(); y = ys[next() % 12]; break;                                   /* utilisations */
      case 1: x = exp((uni() - 0.3) * 40.0); y = (uni() - 0.5) * 20.0; break;          /* wide range */
      case 2: x = (double)(next() % 2000000 + 1); y = ys[next() % 12]; break;          /* integers (slack ** 0.6) */
      case 3: x = uni() * 1000.0 + 1e-9; y = uni() * 4.0; break;
      /* outside the fast path: subnormal x, tiny and huge |y|, results through overflow and underflow */
      case 4: x = dbl_of(next() & 0x000fffffffffffffull | 1ull); y = (uni() - 0.5) * 3.0; break;
      case 5: x = exp((uni() - 0.5) * 8.0); y = ldexp(uni() - 0.5, (int)(next() % 60) - 110); break;
      case 6: x = exp((uni() - 0.5) * 8.0); y = ldexp(uni() - 0.5, 62 + (int)(next() % 30)); break;
      default: x = exp((uni() - 0.5) * 1400.0); y = (uni() - 0.5) * 8.0; break;
    }
    if (y == 0.0 || !(x > 0.0)) continue;
    int ok;
    double got = ks_glibc_pow_fast(x, y, &ok);
    if (!ok) {  /* only the unported band may decline: 512 <= |y log x| < 1024 */
      double e = fabs(y * log(x));
      if (!(e > 500.0 && e < 1030.0)) { if (bad < 10) printf("DECLINED outside the band x=%a y=%a\n", x, y); bad++; }
      continue;
    }
    fast++;
    double want = pow(x, y);
    if (bits_of(got) != bits_of(want)) {
      if (bad < 10) printf("MISMATCH x=%a y=%a got=%a want=%a\n", x, y, got, want);
      bad++;
    }
  }
  printf("pow: checked %ld pairs, %ld on the fast path, %ld mismatches\n", n, fast, bad);
  /* exp and log: the whole domain, no fallback */
  long ebad = 0, lbad = 0;
  for (long it = 0; it < n; ++it) {
    double x;
    switch (it & 7) {
      case 0: x = (uni() - 0.5) * 4.0; break;
      case 1: x = (uni() - 0.5) * 1500.0; break;                 /* through overflow and the subnormal results */
      case 2: x = -700.0 - uni() * 50.0; break;                  /* subnormal results */
      case 3: x = dbl_of(next()); break;                         /* any bit pattern */
      case 4: x = (uni() - 0.5) * 1e-10; break;
      case 5: x = 700.0 + uni() * 12.0; break;
      default: x = (uni() - 0.5) * 60.0; break;
    }
    double got = ks_glibc_exp(x), want = exp(x);
    if (bits_of(got) != bits_of(want) && !(got != got && want != want)) {
      if (ebad < 10) printf("EXP MISMATCH x=%a got=%a want=%a\n", x, got, want);
      ebad++;
    }
    switch (it & 7) {
      case 0: x = 0.9 + uni() * 0.2; break;                      /* the near-1 polynomial and its borders */
      case 1: x = exp((uni() - 0.5) * 1400.0); break;
      case 2: x = dbl_of(next() & 0x7fffffffffffffffull); break; /* any non-negative bit pattern, subnormals included */
      case 3: x = (double)(next() % 2000000 + 1); break;
      case 4: x = dbl_of(next() % 0x10000000000000ull); break;   /* subnormals */
      case 5: x = 1.0 + (uni() - 0.5) * 1e-8; break;
      default: x = uni() * 1000.0; break;
    }
    if (x != x || x == 0.0) continue;
    got = ks_glibc_log(x), want = log(x);
    if (bits_of(got) != bits_of(want)) {
      if (lbad < 10) printf("LOG MISMATCH x=%a got=%a want=%a\n", x, got, want);
      lbad++;
    }
  }
  printf("exp: %ld operands, %ld mismatches; log: %ld mismatches\n", n, ebad, lbad);
  return bad != 0 || ebad != 0 || lbad != 0;
}
