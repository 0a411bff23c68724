// ks_glibc_explog.h -- glibc 2.39's exp() and log() for x86-64 with FMA (__exp_fma, __log_fma), operation by
// operation like ks_glibc_pow.h: which products and sums are fused follows the machine code of that build, not
// the C source (sysdeps/ieee754/dbl-64/e_exp.c, e_log.c), because math.exp / math.log in a policy return exactly
// what that code returns and one ulp can flip int(score).
//
// Included by ks_sim_kernel.cu (device) and tools/check_glibc_pow.c (host, compared with the running libm).
// Needs the KS_PF_* macros and ks_glibc_pow_tables.inc (see ks_glibc_pow.h).
//
// ks_glibc_exp: every double (overflow gives +inf, underflow gives subnormals / 0 as libm does; errno is the caller's
// business).  ks_glibc_log: x > 0 or NaN (the caller raises ValueError for x <= 0 before calling).
KS_PF_FN double ks_glibc_exp(double x) {
  typedef unsigned long long u64;
  const u64 ixb = KS_PF_BITS(x);
  unsigned abstop = (unsigned)(ixb >> 52) & 0x7ffu;
  if (abstop - 0x3c9u > 0x3eu) {
    if ((int)(abstop - 0x3c9u) < 0) return KS_PF_ADD(x, 1.0);  // |x| < 2^-54
    if (abstop > 0x408u) {                                      // |x| >= 1024, inf or nan
      if (ixb == 0xfff0000000000000ull) return 0.0;
      if (abstop == 0x7ffu) return KS_PF_ADD(x, 1.0);
      return (ixb >> 63) ? 0.0 : KS_PF_DBL(0x7ff0000000000000ull);
    }
    abstop = 0;  // 512 <= |x| < 1024: the scale may leave the normal range
  }
  double kd = KS_PF_FMA(x, KS_EXP_INVLN2N, KS_EXP_SHIFT);
  const u64 ki = KS_PF_BITS(kd);
  kd = KS_PF_SUB(kd, KS_EXP_SHIFT);
  const double r = KS_PF_FMA(kd, KS_EXP_NEGLN2LON, KS_PF_FMA(kd, KS_EXP_NEGLN2HIN, x));
  const unsigned idx = 2u * (unsigned)(ki & 127u);
  u64 sbits = ks_exp_tab[idx + 1] + (ki << 45);
  const double tail = KS_PF_DBL(ks_exp_tab[idx]);
  const double p23 = KS_PF_FMA(r, ks_exp_C[1], ks_exp_C[0]);
  const double s = KS_PF_ADD(r, tail);
  const double r2 = KS_PF_MUL(r, r);
  const double p45 = KS_PF_FMA(r, ks_exp_C[3], ks_exp_C[2]);
  const double t = KS_PF_FMA(p23, r2, s);
  const double r4 = KS_PF_MUL(r2, r2);
  const double tmp = KS_PF_FMA(r4, p45, t);
  if (abstop != 0) {
    const double scale = KS_PF_DBL(sbits);
    return KS_PF_FMA(scale, tmp, scale);
  }
  if ((ki & 0x80000000ull) == 0) {  // k > 0: the exponent of scale may have overflowed by <= 460
    sbits -= 1009ull << 52;
    const double scale = KS_PF_DBL(sbits);
    return KS_PF_MUL(KS_PF_FMA(scale, tmp, scale), 0x1p1009);
  }
  sbits += 1022ull << 52;           // k < 0: care in the subnormal range
  const double scale = KS_PF_DBL(sbits);
  const double st = KS_PF_MUL(tmp, scale);
  double y = KS_PF_ADD(scale, st);
  if (y < 1.0) {
    const double hi = KS_PF_ADD(y, 1.0);
    const double lo = KS_PF_ADD(KS_PF_SUB(scale, y), st);
    y = KS_PF_ADD(KS_PF_SUB(1.0, hi), y);
    y = KS_PF_ADD(y, lo);
    y = KS_PF_ADD(y, hi);
    y = KS_PF_SUB(y, 1.0);
    if (y == 0.0) y = 0.0;
  }
  return KS_PF_MUL(y, 0x1p-1022);
}

KS_PF_FN double ks_glibc_log(double x) {
  typedef unsigned long long u64;
  u64 ix = KS_PF_BITS(x);
  if (ix - 0x3fee000000000000ull < 0x3090000000000ull) {  // 1 - 2^-4 <= x < 1 + 0x1.09p-4
    if (ix == 0x3ff0000000000000ull) return 0.0;
    const double r = KS_PF_SUB(x, 1.0);
    const double b12 = KS_PF_FMA(r, ks_log_B[2], ks_log_B[1]);
    const double b45 = KS_PF_FMA(r, ks_log_B[5], ks_log_B[4]);
    const double r2 = KS_PF_MUL(r, r);
    const double b78 = KS_PF_FMA(r, ks_log_B[8], ks_log_B[7]);
    const double b123 = KS_PF_FMA(r2, ks_log_B[3], b12);
    const double b456 = KS_PF_FMA(r2, ks_log_B[6], b45);
    const double r3 = KS_PF_MUL(r, r2);
    double q = KS_PF_FMA(r2, ks_log_B[9], b78);
    q = KS_PF_FMA(r3, ks_log_B[10], q);
    q = KS_PF_FMA(q, r3, b456);
    q = KS_PF_FMA(q, r3, b123);
    const double w = KS_PF_FMA(r, 0x1p27, r);
    const double rhi = KS_PF_FMA(-0x1p27, r, w);
    const double rhi2 = KS_PF_MUL(rhi, rhi);
    const double rlo = KS_PF_SUB(r, rhi);
    const double hi = KS_PF_FMA(rhi2, ks_log_B[0], r);
    const double rmhi = KS_PF_SUB(r, hi);
    const double rsum = KS_PF_ADD(r, rhi);
    double lo = KS_PF_FMA(rhi2, ks_log_B[0], rmhi);
    lo = KS_PF_FMA(KS_PF_MUL(ks_log_B[0], rlo), rsum, lo);
    const double y = KS_PF_FMA(q, r3, lo);
    return KS_PF_ADD(hi, y);
  }
  const unsigned top = (unsigned)(ix >> 48);
  if (top - 0x10u >= 0x7ff0u - 0x10u) {
    if (ix * 2 == 0) return -KS_PF_DBL(0x7ff0000000000000ull);    // log(0) = -inf (the caller raised before)
    if (ix == 0x7ff0000000000000ull) return x;                    // log(inf) = inf
    if ((top & 0x8000u) || (top & 0x7ff0u) == 0x7ff0u) return KS_PF_DBL(0x7ff8000000000000ull);  // negative or NaN
    ix = KS_PF_BITS(KS_PF_MUL(x, 0x1p52));                        // subnormal: normalise
    ix -= 52ull << 52;
  }
  const u64 tmp = ix - 0x3fe6000000000000ull;
  const int i = (int)((tmp >> 45) & 127u);
  const int k = (int)((long long)tmp >> 52);
  const double z = KS_PF_DBL(ix - (tmp & 0xfff0000000000000ull));
  const double kd = KS_PF_I2D(k);
  const double invc = ks_log_tab[2 * i], logc = ks_log_tab[2 * i + 1];
  const double w = KS_PF_FMA(kd, KS_POW_LN2HI, logc);
  const double r = KS_PF_FMA(z, invc, -1.0);
  const double a12 = KS_PF_FMA(r, ks_log_A[2], ks_log_A[1]);
  const double hi = KS_PF_ADD(r, w);
  const double r2 = KS_PF_MUL(r, r);
  double lo = KS_PF_ADD(KS_PF_SUB(w, hi), r);
  lo = KS_PF_FMA(kd, KS_POW_LN2LO, lo);
  const double r3 = KS_PF_MUL(r, r2);
  const double a34 = KS_PF_FMA(r, ks_log_A[4], ks_log_A[3]);
  lo = KS_PF_FMA(r2, ks_log_A[0], lo);
  const double p = KS_PF_FMA(a34, r2, a12);
  const double y = KS_PF_FMA(r3, p, lo);
  return KS_PF_ADD(y, hi);
}
