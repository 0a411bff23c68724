// ks_glibc_pow.h -- the fast path of glibc 2.39's pow() for x86-64 with FMA (__pow_fma), operation by
// operation: which multiplications and additions are fused follows the machine code of that build (GCC
// contracted several a*b+c of the C source, sysdeps/ieee754/dbl-64/e_pow.c, into FMAs), because CPython's
// float ** float returns exactly what that code returns and a one-ulp difference can flip int(score).
//
// Included twice: by ks_sim_kernel.cu (device: __fma_rn/__dadd_rn/__dmul_rn, so nothing is re-fused) and by
// tools/check_glibc_pow.c (host: validated against the running libm on tens of millions of operands).
// The includer defines KS_PF_FMA(a,b,c), KS_PF_ADD, KS_PF_SUB, KS_PF_MUL, KS_PF_BITS(double)->u64,
// KS_PF_DBL(u64)->double, KS_PF_I2D(int)->double, KS_PF_FN (function qualifiers) and the tables of
// ks_glibc_pow_tables.inc.
//
// Domain: x positive and finite (subnormals included), y finite and non-zero.  The function returns exactly what glibc
// returns -- including its shortcuts for |y| < 2^-65 (1.0), |y| >= 2^63 and |y log x| >= 1024 (overflow to +inf, which
// CPython turns into OverflowError, or underflow to 0.0) and |y log x| < 2^-54 (1.0) -- except in the band
// 512 <= |y log x| < 1024 (results beyond 2^+-738), where glibc takes extra care of subnormal results that is not
// ported: there *ok is cleared and the caller reports the operation as outside the translated subset.
KS_PF_FN double ks_glibc_pow_fast(double x, double y, int* ok) {
  typedef unsigned long long u64;
  u64 ix = KS_PF_BITS(x);
  const u64 iy = KS_PF_BITS(y);
  unsigned topx = (unsigned)(ix >> 52);
  const unsigned topy = (unsigned)(iy >> 52);
  *ok = 1;
  if ((topy & 0x7ffu) - 0x3beu > 0x7fu) {  // |y| < 2^-65 or |y| >= 2^63 (e_pow.c: "Special cases" for finite non-zero y)
    if (ix == 0x3ff0000000000000ull) return 1.0;
    if ((topy & 0x7ffu) < 0x3beu) return 1.0;  // 1.0 + y or 1.0 - y, which round to 1.0
    return ((ix > 0x3ff0000000000000ull) == (topy < 0x800u)) ? KS_PF_DBL(0x7ff0000000000000ull) : 0.0;
  }
  if (topx - 1u > 0x7fdu) {
    if (topx != 0u || ix == 0ull) {  // zero, negative, inf, nan: the callers handle these before calling
      *ok = 0;
      return 0.0;
    }
    // subnormal x: normalise (e_pow.c: ix = asuint64(x * 0x1p52); ix &= 0x7fff...; ix -= 52ULL << 52)
    ix = KS_PF_BITS(KS_PF_MUL(x, 4503599627370496.0)) & 0x7fffffffffffffffull;
    ix -= 52ull << 52;
    topx = (unsigned)(ix >> 52);
  }
  // ---- log_inline ----
  const u64 tmp = ix - 0x3fe6955500000000ull;
  const int i = (int)((tmp >> 45) & 127u);
  const int k = (int)((long long)tmp >> 52);
  const double z = KS_PF_DBL(ix - (tmp & 0xfff0000000000000ull));
  const double kd = KS_PF_I2D(k);
  const double invc = ks_pow_log_tab[3 * i], logc = ks_pow_log_tab[3 * i + 1], logctail = ks_pow_log_tab[3 * i + 2];
  const double r = KS_PF_FMA(z, invc, -1.0);
  const double t1 = KS_PF_FMA(kd, KS_POW_LN2HI, logc);
  const double lo1 = KS_PF_FMA(kd, KS_POW_LN2LO, logctail);
  const double ar = KS_PF_MUL(r, ks_pow_A[0]);
  const double q12 = KS_PF_FMA(r, ks_pow_A[2], ks_pow_A[1]);
  const double q34 = KS_PF_FMA(r, ks_pow_A[4], ks_pow_A[3]);
  const double t2 = KS_PF_ADD(r, t1);
  const double lo2 = KS_PF_ADD(KS_PF_SUB(t1, t2), r);
  const double ar2 = KS_PF_MUL(r, ar);
  const double ar3 = KS_PF_MUL(r, ar2);
  const double lo3 = KS_PF_FMA(ar, r, -ar2);
  const double hi = KS_PF_ADD(t2, ar2);
  const double q56 = KS_PF_FMA(r, ks_pow_A[6], ks_pow_A[5]);
  const double lo4 = KS_PF_ADD(KS_PF_SUB(t2, hi), ar2);
  const double q3456 = KS_PF_FMA(q56, ar2, q34);
  const double pin = KS_PF_FMA(ar2, q3456, q12);
  const double s3 = KS_PF_ADD(KS_PF_ADD(KS_PF_ADD(lo1, lo2), lo3), lo4);
  const double lo = KS_PF_FMA(ar3, pin, s3);
  const double lhi = KS_PF_ADD(hi, lo);
  const double ltail = KS_PF_ADD(KS_PF_SUB(hi, lhi), lo);
  // ---- y * log(x) in two pieces ----
  const double ehi = KS_PF_MUL(y, lhi);
  const double elo = KS_PF_FMA(y, ltail, KS_PF_FMA(lhi, y, -ehi));
  // ---- exp_inline ----
  const unsigned abstop = (unsigned)(KS_PF_BITS(ehi) >> 52) & 0x7ffu;
  if (abstop - 0x3c9u > 0x3eu) {
    if (abstop < 0x3c9u) return 1.0;  // |y log x| < 2^-54: 1.0 + ehi rounds to 1.0
    if (abstop >= 0x409u) return (KS_PF_BITS(ehi) >> 63) ? 0.0 : KS_PF_DBL(0x7ff0000000000000ull);  // >= 1024: underflow / overflow
    *ok = 0;  // 512 <= |y log x| < 1024: glibc's specialcase() path, not ported
    return 0.0;
  }
  double kd2 = KS_PF_FMA(ehi, KS_EXP_INVLN2N, KS_EXP_SHIFT);
  const u64 ki = KS_PF_BITS(kd2);
  kd2 = KS_PF_SUB(kd2, KS_EXP_SHIFT);
  double rr = KS_PF_FMA(kd2, KS_EXP_NEGLN2LON, KS_PF_FMA(kd2, KS_EXP_NEGLN2HIN, ehi));
  rr = KS_PF_ADD(elo, rr);
  const unsigned idx = 2u * (unsigned)(ki & 127u);
  const u64 sbits = ks_exp_tab[idx + 1] + (ki << 45);
  const double etail = KS_PF_DBL(ks_exp_tab[idx]);
  const double p23 = KS_PF_FMA(rr, ks_exp_C[1], ks_exp_C[0]);
  const double s = KS_PF_ADD(rr, etail);
  const double r2 = KS_PF_MUL(rr, rr);
  const double p45 = KS_PF_FMA(rr, ks_exp_C[3], ks_exp_C[2]);
  const double t = KS_PF_FMA(p23, r2, s);
  const double r4 = KS_PF_MUL(r2, r2);
  const double tmp2 = KS_PF_FMA(p45, r4, t);
  const double scale = KS_PF_DBL(sbits);
  return KS_PF_FMA(tmp2, scale, scale);
}
