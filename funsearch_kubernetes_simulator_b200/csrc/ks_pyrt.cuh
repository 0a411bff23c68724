// ks_pyrt.cuh -- Python-number semantics for NVRTC-compiled policy functions.
//
// The reference runs policies as CPython code (funsearch/funsearch_integration.py:77-101,
// restricted environment funsearch/safe_execution.py:98-124).  Everything a translated
// policy computes must therefore follow CPython 3.12 arithmetic:
// true division, floor // and %, float_pow special cases, int() truncation with
// OverflowError/ValueError, max/min returning the first extremal operand, the
// Neumaier-compensated builtin sum(), exceptions aborting the episode.
//
// Conventions for generated code
//   ks_int  = Python int restricted to int64; leaving int64 => KS_RET_UNSUPPORTED (never a wrap)
//   double  = Python float          int (0/1) = Python bool
//   KsNum   = a value whose int-or-float type is only known at run time
//   every helper that can raise takes `int* st`; st==1 => exception, st==2 => unsupported.
// Compile with --fmad=false: CPython never fuses a*b+c.
#ifndef KS_PYRT_CUH_
#define KS_PYRT_CUH_

#include "ks_abi.h"

typedef long long ks_int;
typedef unsigned long long ks_u64;

#define KS_D __device__ __forceinline__

// The status word is sticky and remembers the FIRST failure (1 = exception, 2 = unsupported): helpers
// never trap or fault on bad operands, so a translated policy keeps computing and tests the word only
// where it returns.  Any other outcome after a failure is unobservable: the episode is over.
#define KS_RETURN(x) return (st ? (st == 1 ? KS_RET_EXCEPTION : KS_RET_UNSUPPORTED) : (long long)(x))
__device__ __forceinline__ void ks_fail(int* st, int code) {
  if (!*st) *st = code;
}

// Heavy libm-class routines live once in the kernel module (ks_sim_kernel.cu) and are linked in.
extern "C" __device__ double ks_rt_pow(double x, double y);
extern "C" __device__ double ks_rt_exp(double x);
extern "C" __device__ double ks_rt_log(double x);
extern "C" __device__ ks_int ks_rt_idiv(ks_int a, ks_int b);  // C truncating quotient, b != 0
// `return max(lo, int(score))` as the scheduler wrapper sees it: int(max(0, .)) or the exception channel
extern "C" __device__ long long ks_rt_ret_max_f2i(double score, long long lo, int st);

// ------------------------------------------------------------------------------- field access
// mutable node state: a plain shared-memory load (not volatile: the state does not change while a policy
// is being evaluated, so the compiler may merge repeated reads of one field)
KS_D int ks_ns(unsigned ns, int field) {
  int v;
  asm("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(ns + 4u * (unsigned)field));
  return v;
}
KS_D int ks_nc(const int* nc, int field) { return __ldg(nc + field); }
// float literal i of the policy being evaluated (the structure's code is shared by all policies that differ only in
// these values); a warp-uniform read-only load
KS_D double ks_kf(const double* kf, int i) { return __ldg(kf + i); }
// fp64 division, shared by every policy of the module (keeps the per-policy code small)
extern "C" __device__ double ks_rt_ddiv(double a, double b);
#ifdef KS_INLINE_DIV
#define ks_div(a, b) __ddiv_rn((a), (b))
#else
#define ks_div(a, b) ks_rt_ddiv((a), (b))
#endif

// ------------------------------------------------------------------------------- int64 with overflow
KS_D ks_int ks_iadd(ks_int a, ks_int b, int* st) {
  ks_int r = (ks_int)((ks_u64)a + (ks_u64)b);
  if (((a ^ r) & (b ^ r)) < 0) ks_fail(st, 2);
  return r;
}
KS_D ks_int ks_isub(ks_int a, ks_int b, int* st) {
  ks_int r = (ks_int)((ks_u64)a - (ks_u64)b);
  if (((a ^ b) & (a ^ r)) < 0) ks_fail(st, 2);
  return r;
}
KS_D ks_int ks_imul(ks_int a, ks_int b, int* st) {
  ks_int lo = (ks_int)((ks_u64)a * (ks_u64)b);
  ks_int hi = __mul64hi(a, b);
  if (hi != (lo >> 63)) ks_fail(st, 2);
  return lo;
}
KS_D ks_int ks_ineg(ks_int a, int* st) {
  if (a == (ks_int)0x8000000000000000LL) ks_fail(st, 2);
  return -a;
}
KS_D ks_int ks_iabs(ks_int a, int* st) { return a < 0 ? ks_ineg(a, st) : a; }
// Python floor division / modulo on ints: sign of the divisor, ZeroDivisionError on 0.
KS_D ks_int ks_ifloordiv(ks_int a, ks_int b, int* st) {
  if (b == 0) { ks_fail(st, 1); return 0; }
  if (b == -1) return ks_ineg(a, st);
  ks_int q = ks_rt_idiv(a, b);
  ks_int r = a - q * b;
  if (r != 0 && ((r < 0) != (b < 0))) q -= 1;
  return q;
}
KS_D ks_int ks_imod(ks_int a, ks_int b, int* st) {
  if (b == 0) { ks_fail(st, 1); return 0; }
  if (b == -1) return 0;
  ks_int q = ks_rt_idiv(a, b);
  ks_int r = a - q * b;
  if (r != 0 && ((r < 0) != (b < 0))) r += b;
  return r;
}
// int -> float as CPython does it (round to nearest even; exact below 2**53).
KS_D double ks_i2d(ks_int a) { return __ll2double_rn(a); }
// int / int: CPython rounds the exact quotient once.  Below 2**53 both operands convert exactly and
// the IEEE division is that single rounding; larger operands would need the long-division path.
KS_D double ks_itruediv(ks_int a, ks_int b, int* st) {
  if (b == 0) { ks_fail(st, 1); return 0.0; }
  const ks_int lim = 1LL << 53;
  if (a > lim || a < -lim || b > lim || b < -lim) { ks_fail(st, 2); return 0.0; }
  return ks_div((double)a, (double)b);
}
// int ** non-negative int, exact.
KS_D ks_int ks_ipow(ks_int base, ks_int e, int* st) {
  ks_int r = 1;
  if (base == 0 || base == 1) return e == 0 ? 1 : base;
  if (base == -1) return (e & 1) ? -1 : 1;
  if (e > 63) { ks_fail(st, 2); return 0; }
  while (e > 0) {
    if (e & 1) r = ks_imul(r, base, st);
    e >>= 1;
    if (e) base = ks_imul(base, base, st);
    if (*st) return 0;
  }
  return r;
}

// ------------------------------------------------------------------------------- float
KS_D double ks_ftruediv(double a, double b, int* st) {
  if (b == 0.0) { ks_fail(st, 1); return 0.0; }
  return ks_div(a, b);
}
// float % float (Objects/floatobject.c float_rem)
KS_D double ks_fmod(double a, double b, int* st) {
  if (b == 0.0) { ks_fail(st, 1); return 0.0; }
  double m = fmod(a, b);
  if (m != 0.0) {
    if ((b < 0.0) != (m < 0.0)) m += b;
  } else {
    m = copysign(0.0, b);
  }
  return m;
}
// float // float (Objects/floatobject.c _float_div_mod + float_floor_div)
KS_D double ks_ffloordiv(double a, double b, int* st) {
  if (b == 0.0) { ks_fail(st, 1); return 0.0; }
  double m = fmod(a, b);
  double div = ks_div(a - m, b);
  if (m != 0.0) {
    if ((b < 0.0) != (m < 0.0)) div -= 1.0;
  }
  double fl;
  if (div != 0.0) {
    fl = floor(div);
    if (div - fl > 0.5) fl += 1.0;
  } else {
    fl = copysign(0.0, ks_div(a, b));
  }
  return fl;
}
// float ** float and math.pow live once in the kernel module (csrc/ks_sim_kernel.cu): they return the
// result, or a NaN with a reserved payload when CPython would raise (1) / the result is complex (2).
#define KS_NAN_TAG 0x7ff8dead
extern "C" __device__ double ks_rt_fpow(double x, double y);
extern "C" __device__ double ks_rt_mpow(double x, double y);
KS_D double ks_untag(double r, int* st) {
  const long long b = __double_as_longlong(r);
  if ((int)(b >> 32) == KS_NAN_TAG) {
    ks_fail(st, (int)(b & 3));
    return 0.0;
  }
  return r;
}
KS_D double ks_fpow(double x, double y, int* st) { return ks_untag(ks_rt_fpow(x, y), st); }
// int ** negative int goes through float pow (Objects/longobject.c long_pow -> float_pow)
KS_D double ks_ipow_neg(ks_int a, ks_int b, int* st) { return ks_fpow(ks_i2d(a), ks_i2d(b), st); }

// int(x) for float x: truncation; OverflowError on inf, ValueError on nan.
KS_D ks_int ks_f2i(double x, int* st) {
  if (!(fabs(x) < 9223372036854775808.0)) {  // nan, inf, or beyond int64 (one test on the common path)
    ks_fail(st, (isnan(x) || isinf(x)) ? 1 : 2);
    return 0;
  }
  return (ks_int)x;  // cvt.rzi
}
// round(x) with one argument: half to even, result is an int.
KS_D ks_int ks_fround(double x, int* st) {
  const double r = rint(x);
  if (!(fabs(r) < 9223372036854775808.0)) {
    ks_fail(st, (isnan(x) || isinf(x)) ? 1 : 2);
    return 0;
  }
  return (ks_int)r;
}
// round(x, ndigits) for float x: CPython (Objects/floatobject.c double_round) prints x with ndigits decimals correctly
// rounded (half to even on the EXACT binary value) and reads the decimal string back, i.e. the result is the double
// nearest to R / 10^ndigits where R = round_half_even(x * 10^ndigits) computed without error.  Here: x = m * 2^e exactly,
// m * 10^ndigits in 128 bits, shifted right by -e with the tie rule; R / 10^ndigits is one correctly rounded IEEE
// division because both are exact doubles (R < 2^53, ndigits <= 15).  Outside that range: 'unsupported', never a guess.
KS_D double ks_fround_n(double x, ks_int nd, int* st) {
  if (isnan(x) || isinf(x) || x == 0.0) return x;
  if (nd < 0 || nd > 15) { ks_fail(st, 2); return 0.0; }
  const double ax = fabs(x);
  if (ax >= 4503599627370496.0) return x;  // 2^52: already an integer
  const ks_u64 bits = (ks_u64)__double_as_longlong(ax);
  int e = (int)(bits >> 52);
  ks_u64 m = bits & 0xfffffffffffffull;
  if (e == 0) e = 1; else m |= 1ull << 52;
  e -= 1075;  // ax = m * 2^e, e < 0 here
  ks_u64 p10 = 1;
  for (ks_int k = 0; k < nd; ++k) p10 *= 10ull;  // <= 10^15 < 2^50
  const ks_u64 lo = m * p10, hi = __umul64hi(m, p10);  // m * 10^nd < 2^103
  const int sh = -e;  // 1 .. 1074: R = (m * 10^nd) >> sh, rounded half to even on the bits shifted out
  ks_u64 q = 0, rem_hi = 0, rem_lo = 0, half_hi = 1, half_lo = 0;  // defaults: remainder 0 < half -> rounds down to q
  if (sh > 103) {
    q = 0;  // the product is below 2^103 <= 2^(sh-1): less than one half, rounds to zero
  } else if (sh < 64) {
    if (hi >> sh) { ks_fail(st, 2); return 0.0; }  // R would not fit 64 bits
    q = (lo >> sh) | (hi << (64 - sh));  // sh >= 1, so the left shift is by at most 63
    rem_lo = lo & ((1ull << sh) - 1ull);
    half_hi = 0;
    half_lo = 1ull << (sh - 1);
  } else if (sh == 64) {
    q = hi;
    rem_lo = lo;
    half_hi = 0;
    half_lo = 1ull << 63;
  } else {
    const int s2 = sh - 64;  // 1 .. 39
    q = hi >> s2;
    rem_hi = hi & ((1ull << s2) - 1ull);
    rem_lo = lo;
    half_hi = 1ull << (s2 - 1);
    half_lo = 0;
  }
  const bool above = rem_hi > half_hi || (rem_hi == half_hi && rem_lo > half_lo);
  const bool tie = rem_hi == half_hi && rem_lo == half_lo;
  if (above || (tie && (q & 1ull))) q += 1;
  if (q >= (1ull << 53)) { ks_fail(st, 2); return 0.0; }
  const double r = ks_div((double)q, (double)p10);
  return x < 0.0 ? -r : r;
}
KS_D double ks_msqrt(double x, int* st) {
  if (x < 0.0) { ks_fail(st, 1); return 0.0; }  // ValueError: math domain error
  return __dsqrt_rn(x);
}
KS_D double ks_mlog(double x, int* st) {
  if (isnan(x)) return x;
  if (x <= 0.0) { ks_fail(st, 1); return 0.0; }
  return ks_rt_log(x);
}
KS_D double ks_mexp(double x, int* st) {
  if (isnan(x)) return x;
  double r = ks_rt_exp(x);
  if (isinf(r) && !isinf(x)) { ks_fail(st, 1); return 0.0; }  // OverflowError: math range error
  return r;
}
// math.sin / cos / tan: CPython returns glibc's results, which are not correctly rounded, and this runtime has no
// operation-for-operation port of them (as it has for pow, exp and log).  A last-place difference can flip int(score),
// so under the bit-exact contract they are outside the translated subset (status 'unsupported'), never approximated.
// The domain error is still an exception: math.sin(inf) raises ValueError whatever the library.
KS_D double ks_msin(double x, int* st) {
  ks_fail(st, isinf(x) ? 1 : 2);
  return 0.0;
}
KS_D double ks_mcos(double x, int* st) { return ks_msin(x, st); }
KS_D double ks_mtan(double x, int* st) { return ks_msin(x, st); }
KS_D double ks_mpow(double x, double y, int* st) { return ks_untag(ks_rt_mpow(x, y), st); }

// ------------------------------------------------------------------------------- exact int <-> float order
// Python compares an int with a float by value, never by converting the int with rounding.
KS_D int ks_cmp_if(ks_int i, double f) {  // -1, 0, +1, or 2 when f is nan (all comparisons false)
  if (isnan(f)) return 2;
  if (f >= 9223372036854775808.0) return -1;
  if (f < -9223372036854775808.0) return 1;
  double fl = floor(f);
  ks_int fi = (ks_int)fl;  // exact: |fl| < 2**63 and integral
  if (i < fi) return -1;
  if (i > fi) return 1;
  return (f > fl) ? -1 : 0;
}
KS_D int ks_cmp_ff(double a, double b) {
  if (isnan(a) || isnan(b)) return 2;
  return a < b ? -1 : (a > b ? 1 : 0);
}
KS_D int ks_cmp_ii(ks_int a, ks_int b) { return a < b ? -1 : (a > b ? 1 : 0); }

// ------------------------------------------------------------------------------- dynamically typed numbers
struct KsNum {
  ks_int i;
  double d;
  int f;  // 0 = int (or bool), 1 = float
};
KS_D KsNum ks_num_i(ks_int v) { KsNum n; n.i = v; n.d = 0.0; n.f = 0; return n; }
KS_D KsNum ks_num_f(double v) { KsNum n; n.i = 0; n.d = v; n.f = 1; return n; }
KS_D double ks_num_d(KsNum a) { return a.f ? a.d : ks_i2d(a.i); }
KS_D int ks_num_truth(KsNum a) { return a.f ? (a.d != 0.0) : (a.i != 0); }
KS_D int ks_cmp_nn(KsNum a, KsNum b) {
  if (a.f) {
    if (b.f) return ks_cmp_ff(a.d, b.d);
    int c = ks_cmp_if(b.i, a.d);
    return c == 2 ? 2 : -c;
  }
  if (b.f) return ks_cmp_if(a.i, b.d);
  return ks_cmp_ii(a.i, b.i);
}
KS_D KsNum ks_nadd(KsNum a, KsNum b, int* st) {
  if (!a.f && !b.f) return ks_num_i(ks_iadd(a.i, b.i, st));
  return ks_num_f(ks_num_d(a) + ks_num_d(b));
}
KS_D KsNum ks_nsub(KsNum a, KsNum b, int* st) {
  if (!a.f && !b.f) return ks_num_i(ks_isub(a.i, b.i, st));
  return ks_num_f(ks_num_d(a) - ks_num_d(b));
}
KS_D KsNum ks_nmul(KsNum a, KsNum b, int* st) {
  if (!a.f && !b.f) return ks_num_i(ks_imul(a.i, b.i, st));
  return ks_num_f(ks_num_d(a) * ks_num_d(b));
}
KS_D double ks_ntruediv(KsNum a, KsNum b, int* st) {
  if (!a.f && !b.f) return ks_itruediv(a.i, b.i, st);
  return ks_ftruediv(ks_num_d(a), ks_num_d(b), st);
}
KS_D KsNum ks_nfloordiv(KsNum a, KsNum b, int* st) {
  if (!a.f && !b.f) return ks_num_i(ks_ifloordiv(a.i, b.i, st));
  return ks_num_f(ks_ffloordiv(ks_num_d(a), ks_num_d(b), st));
}
KS_D KsNum ks_nmod(KsNum a, KsNum b, int* st) {
  if (!a.f && !b.f) return ks_num_i(ks_imod(a.i, b.i, st));
  return ks_num_f(ks_fmod(ks_num_d(a), ks_num_d(b), st));
}
KS_D KsNum ks_npow(KsNum a, KsNum b, int* st) {
  if (!a.f && !b.f) {
    if (b.i >= 0) return ks_num_i(ks_ipow(a.i, b.i, st));
    return ks_num_f(ks_ipow_neg(a.i, b.i, st));
  }
  return ks_num_f(ks_fpow(ks_num_d(a), ks_num_d(b), st));
}
KS_D KsNum ks_nneg(KsNum a, int* st) { return a.f ? ks_num_f(-a.d) : ks_num_i(ks_ineg(a.i, st)); }
KS_D KsNum ks_nabs(KsNum a, int* st) { return a.f ? ks_num_f(fabs(a.d)) : ks_num_i(ks_iabs(a.i, st)); }
KS_D ks_int ks_n2i(KsNum a, int* st) { return a.f ? ks_f2i(a.d, st) : a.i; }
KS_D ks_int ks_nround(KsNum a, int* st) { return a.f ? ks_fround(a.d, st) : a.i; }

// ------------------------------------------------------------------------------- builtin sum()
// Python/bltinmodule.c builtin_sum_impl (3.12): ints accumulate exactly; the first float switches to
// a double accumulator; float items are then added with Neumaier compensation, int items plainly.
struct KsSum {
  ks_int i;
  double f;
  double c;
  int mode;  // 0 = int, 1 = float
};
KS_D void ks_sum_init(KsSum* s) { s->i = 0; s->f = 0.0; s->c = 0.0; s->mode = 0; }
KS_D void ks_sum_add_i(KsSum* s, ks_int v, int* st) {
  if (s->mode == 0) s->i = ks_iadd(s->i, v, st);
  else s->f += ks_i2d(v);
}
KS_D void ks_sum_add_f(KsSum* s, double x) {
  if (s->mode == 0) {
    s->f = ks_i2d(s->i) + x;  // PyNumber_Add(int, float)
    s->c = 0.0;
    s->mode = 1;
    return;
  }
  double t = s->f + x;
  if (fabs(s->f) >= fabs(x)) s->c += (s->f - t) + x;
  else s->c += (x - t) + s->f;
  s->f = t;
}
KS_D void ks_sum_add_n(KsSum* s, KsNum v, int* st) {
  if (v.f) ks_sum_add_f(s, v.d);
  else ks_sum_add_i(s, v.i, st);
}
KS_D KsNum ks_sum_result(const KsSum* s) {
  if (s->mode == 0) return ks_num_i(s->i);
  double r = s->f;
  if (s->c != 0.0 && !isinf(s->c) && !isnan(s->c)) r += s->c;
  return ks_num_f(r);
}

// ------------------------------------------------------------------------------- small lists
// `[g for g in node.gpus if ...]`, sorted(...), slices: a list of GPU indices of one node.
struct KsGpuList {
  int n;
  signed char g[KS_MAX_GPUS_PER_NODE];
};
#define KS_LIST_CAP 32
struct KsIntList {
  int n;
  ks_int v[KS_LIST_CAP];
};
struct KsFloatList {
  int n;
  double v[KS_LIST_CAP];
};
// Python slice bounds for lst[a:b] with step 1.
KS_D int ks_slice_bound(ks_int b, int n) {
  if (b < 0) { b += n; if (b < 0) b = 0; }
  if (b > n) b = n;
  return (int)b;
}

// ------------------------------------------------------------------------------- the scheduler wrapper
// FunSearchScheduler.__call__: int(max(0, score))  (funsearch_integration.py:96)
KS_D long long ks_wrap_i(ks_int s) { return s > 0 ? s : 0; }
KS_D long long ks_wrap_f(double s) {
  if (!(s > 0.0)) return 0;  // max(0, nan) -> 0, max(0, -x) -> 0, max(0, 0.0) -> 0
  if (!(s < 9223372036854775808.0)) return isinf(s) ? KS_RET_EXCEPTION : KS_RET_UNSUPPORTED;  // int(inf): OverflowError
  return (long long)s;
}
KS_D long long ks_wrap_n(KsNum s) { return s.f ? ks_wrap_f(s.d) : ks_wrap_i(s.i); }

#endif  // KS_PYRT_CUH_
