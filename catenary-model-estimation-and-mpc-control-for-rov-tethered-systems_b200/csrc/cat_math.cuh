// cat_math.cuh -- scalar device math shared by every kernel of libcatenary_b200 (sm_100a).
//
// CatMath<double>: IEEE fp64 (1e-9 parity mode).  CatMath<float>: fp32 fast mode (1e-4) on the
// FFMA pipe + MUFU (sin.approx / ex2.approx / lg2.approx / rcp / rsqrt).
// Reference formulas are cited per function (paths relative to the reference checkout).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

// ------------------------------------------------------------------------------------------
// Cold-path counters (bench.py "sensitivity"): how many warp-level entries the data-dependent slow branches of
// the rollout take -- [0] the catenary solve's iteration loop (a lane moved beyond the warm step's range or did
// not converge), [1] the full-range sine (|x| >= 1e5, inf, NaN).  Counting is off unless cat_dbg_on is set
// (cat_mpc_stats), and the test sits inside the cold branches, so the common path carries nothing.  One copy
// per translation unit (no relocatable device code); the unit that instantiates the rollout reads its own.
// ------------------------------------------------------------------------------------------
__device__ unsigned long long cat_dbg_counters[4];
__device__ int cat_dbg_on;
__device__ __forceinline__ void cat_dbg_count(int which) {
  if (cat_dbg_on) {
    const unsigned m = __activemask();
    if ((threadIdx.x & 31) == (unsigned)(__ffs(m) - 1)) atomicAdd(&cat_dbg_counters[which], 1ULL);
  }
}

// ------------------------------------------------------------------------------------------
// fp64 elementary functions written for this path.  CUDA's sin()/exp() re-materialise their
// 64-bit literals with UMOV pairs, fetch a coefficient table with LDG and carry inf/nan/huge-
// argument select chains; in the rollout kernel those were 28% of the issue slots (profiles/r1_*).
// Here every coefficient lives in __constant__ memory, so it is a direct c[bank][off] operand
// of the DFMA, and the rare cases take one cold branch to the CUDA routine.
// ------------------------------------------------------------------------------------------
// sin r = r + r z p(z), z = r^2, |r| <= pi/2.  p = degree-7 interpolant of (sin(sqrt z)/sqrt z - 1)/z at the
// Chebyshev nodes of [0, (pi/2)^2] (mpmath, 60 digits): |error| <= 3.6e-17 with the rounded coefficients, two
// terms fewer than the Taylor series needs for the same range.
__constant__ __align__(16) double CAT_SIN_C[8] = {
    -0.16666666666666666,
    0.008333333333333316,
    -0.00019841269841254974,
    2.7557319219163234e-06,
    -2.5052107616996182e-08,
    1.6058977312464087e-10,
    -7.643970296798572e-13,
    2.7314447669863995e-15};
__constant__ double CAT_PI_C[5] = {
    0.31830988618379067154,      // 1/pi
    6755399441055744.0,          // 1.5 * 2^52: fma(x, 1/pi, magic) rounds x/pi to an integer in the low word
    3.141592653589793116,        // pi, first 53 bits
    1.2246467991473532072e-16,   // next 53 bits
    -2.9947698097183396659e-33}; // next
__constant__ double CAT_HYP_S[7] = {    // sinh(r)/r - 1 = z (s0 + z (s1 + ...))
    1.0 / 6.0, 1.0 / 120.0, 1.0 / 5040.0, 1.0 / 362880.0, 1.0 / 39916800.0, 1.0 / 6227020800.0,
    1.0 / 1307674368000.0};
__constant__ double CAT_HYP_C[7] = {    // cosh(r) - 1 = z (c0 + z (c1 + ...))
    1.0 / 2.0, 1.0 / 24.0, 1.0 / 720.0, 1.0 / 40320.0, 1.0 / 3628800.0, 1.0 / 479001600.0,
    1.0 / 87178291200.0};
__constant__ double CAT_LN2_C[4] = {
    1.4426950408889634074,       // 1/ln2
    6755399441055744.0,
    0.69314718055994528623,      // ln2 hi
    2.3190468138462995584e-17};  // ln2 lo

// sin of N independent arguments at once, |error| <~ 2 ulp for |x| < 1e5 (the scaled features live in
// |x| < 1e3); beyond that, and for inf/nan, ONE cold branch per batch to the CUDA routine.  The N Horner
// chains are interleaved, so each coefficient is fetched once per batch and the DFMA latency of one chain
// hides behind the others.
// full-range sine of one argument, out of line: the rollout reaches it only for |x| >= 1e5, inf or NaN, and four
// inlined copies of the library routine's range logic made up a third of the loop's code
static __device__ __noinline__ double cat_sin_cold_f64(double x) { return sin(x); }

template <int N>
__device__ __forceinline__ void cat_sin_vec_f64(double (&a)[N]) {
  double r[N], z[N], p[N];
  int q[N];
  bool bad = false;
#pragma unroll
  for (int n = 0; n < N; ++n) {
    const double t = fma(a[n], CAT_PI_C[0], CAT_PI_C[1]);
    q[n] = __double2loint(t);
    const double qf = t - CAT_PI_C[1];
    // two-term Cody-Waite: |q| < 3.2e4 on the fast range, so the third term (q * 3e-33) is below 1e-28
    r[n] = fma(-qf, CAT_PI_C[3], fma(-qf, CAT_PI_C[2], a[n]));
    z[n] = r[n] * r[n];
    p[n] = CAT_SIN_C[7];
    // |a| >= 1e5, inf or NaN, tested on the high word (integer pipe; 0x40F86A00 = high word of 1e5)
    bad = bad || (__double2hiint(a[n]) & 0x7fffffff) >= 0x40F86A00;
  }
#pragma unroll
  for (int i = 6; i >= 0; --i) {
#pragma unroll
    for (int n = 0; n < N; ++n) p[n] = fma(p[n], z[n], CAT_SIN_C[i]);
  }
#pragma unroll
  for (int n = 0; n < N; ++n) {
    const double v = fma(r[n] * z[n], p[n], r[n]);
    // sin(x) = (-1)^q sin(r)
    r[n] = __hiloint2double(__double2hiint(v) ^ (q[n] << 31), __double2loint(v));
  }
  if (bad) {
    cat_dbg_count(1);
#pragma unroll
    for (int n = 0; n < N; ++n)
      if ((__double2hiint(a[n]) & 0x7fffffff) >= 0x40F86A00) r[n] = cat_sin_cold_f64(a[n]);
  }
#pragma unroll
  for (int n = 0; n < N; ++n) a[n] = r[n];
}

__device__ __forceinline__ double cat_sin_f64(double x) {
  double a[1] = {x};
  cat_sin_vec_f64<1>(a);
  return a[0];
}

// sqrt(x) without the IEEE special-case path and without a branch: MUFU.RSQ64H seed + two coupled Newton
// steps (~1 ulp) for normal positive x.  The special inputs are patched by integer selects on the high word,
// so the routine stays inside the caller's basic block and adds nothing to the FP64 pipe:
//   +-0 and denormals -> x itself (|error| < 1.5e-154 for a denormal), +inf and NaN -> x, negative -> NaN.
__device__ __forceinline__ double cat_sqrt_f64(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double g = x * y;            // ~ sqrt(x)
  double h = 0.5 * y;          // ~ 1 / (2 sqrt(x))
  double e = fma(-g, h, 0.5);
  g = fma(g, e, g);
  h = fma(h, e, h);
  e = fma(-g, g, x);           // residual x - g^2
  g = fma(e, h, g);
  const int hx = __double2hiint(x);
  const bool special = (unsigned)(hx - 0x00100000) >= 0x7fe00000u;   // exponent 0, exponent 0x7ff, or sign set
  const bool negative = (unsigned)hx >= 0x80100000u;                 // below -2^-1022 (incl. -inf, -NaN)
  const double fix = __hiloint2double(negative ? 0x7ff80000 : hx, __double2loint(x));
  return special ? fix : g;
}

// |x| > c ? copysign(c, x) : x  (c > 0; NaN stays NaN like np.clip): one compare and integer selects instead of
// the two NaN-quieting min/max sequences
__device__ __forceinline__ double cat_clamp_sym_f64(double x, double c) {
  const bool big = fabs(x) > c;
  const int hx = __double2hiint(x);
  return __hiloint2double(big ? (__double2hiint(c) | (hx & (int)0x80000000)) : hx,
                          big ? __double2loint(c) : __double2loint(x));
}

// For u >= 0: sh = sinh u, cm = cosh u - 1, sm = sinh u - u, all without cancellation at small u and
// without a division: u = k ln2 + r, |r| <= ln2/2;  sinh/cosh(k ln2) are exact combinations of
// 2^k, 2^-k;  addition theorems.  For k = 0 this IS the series (A = 0, B = 1).
__device__ __forceinline__ void cat_hyp_f64(double u, double& sm, double& cm, double& sh) {
  u = (u < 700.0) ? u : 700.0;
  const double t = fma(u, CAT_LN2_C[0], CAT_LN2_C[1]);
  const int k = __double2loint(t);
  const double kf = t - CAT_LN2_C[1];
  const double r = fma(-kf, CAT_LN2_C[3], fma(-kf, CAT_LN2_C[2], u));
  const double z = r * r;
  double S = CAT_HYP_S[6], Cc = CAT_HYP_C[6];
#pragma unroll
  for (int i = 5; i >= 0; --i) {
    S = fma(S, z, CAT_HYP_S[i]);
    Cc = fma(Cc, z, CAT_HYP_C[i]);
  }
  const double rS = r * (z * S);          // sinh r - r
  const double chr1 = z * Cc;             // cosh r - 1
  const double shr = rS + r;              // sinh r
  // 2^(k-1), 2^(-k-1)
  const double p2 = __hiloint2double((k + 1022) << 20, 0);
  const double m2 = __hiloint2double((1022 - k) << 20, 0);
  const double A = p2 - m2;               // sinh(k ln2)   (0 for k = 0)
  const double Bm1 = (p2 - 0.5) + m2 - 0.5;   // cosh(k ln2) - 1 (exactly 0 for k = 0)
  // sinh u = A cosh r + B sinh r ; cosh u - 1 = (B - 1) + B (cosh r - 1) + A sinh r
  const double Ach = fma(A, chr1, A);
  sh = fma(Bm1, shr, shr) + Ach;
  cm = fma(Bm1, chr1, chr1) + fma(A, shr, Bm1);
  // sinh u - u = [A cosh r - k ln2] + (B - 1) sinh r + (sinh r - r)
  sm = fma(Bm1, shr, rS) + (fma(-kf, CAT_LN2_C[2], Ach) - kf * CAT_LN2_C[3]);
}

template <typename T> struct CatMath;

template <> struct CatMath<double> {
  typedef double real;
  static __device__ __forceinline__ double sin_(double x) { return cat_sin_f64(x); }
  static __device__ __forceinline__ double cos_(double x) { return cos(x); }
  template <int N> static __device__ __forceinline__ void sin_v(double (&a)[N]) { cat_sin_vec_f64<N>(a); }
  // max(0, x) as np.maximum(0.0, x) (NaN stays NaN), decided on the sign bit: two integer selects instead of
  // the DSETP.MAX + NaN-quieting sequence the compiler emits for x > 0 ? x : 0
  static __device__ __forceinline__ double pos_(double x) {
    const int hi = __double2hiint(x);
    return __hiloint2double(hi < 0 ? 0 : hi, hi < 0 ? 0 : __double2loint(x));
  }
  static __device__ __forceinline__ void sincos_(double x, double* s, double* c) { sincos(x, s, c); }
  static __device__ __forceinline__ double tanh_(double x) { return tanh(x); }
  static __device__ __forceinline__ double exp_(double x) { return exp(x); }
  static __device__ __forceinline__ double log_(double x) { return log(x); }
  static __device__ __forceinline__ double sqrt_(double x) { return cat_sqrt_f64(x); }
  static __device__ __forceinline__ double abs_(double x) { return fabs(x); }
  static __device__ __forceinline__ double rcp_(double x) { return 1.0 / x; }
  static __device__ __forceinline__ double div_(double a, double b) { return a / b; }
  // 1/x for normal positive x to ~1 ulp without the IEEE fix-up path: MUFU.RCP64H seed (20 bits) + two
  // Newton refinements.  Used where the quotient is a correction or feeds a 1e-9 comparison.
  static __device__ __forceinline__ double rcp_fast_(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    return fma(y, e, y);
  }
  static __device__ __forceinline__ double clamp_sym_(double x, double c) { return cat_clamp_sym_f64(x, c); }
  static __device__ __forceinline__ double min_(double a, double b) { return fmin(a, b); }
  static __device__ __forceinline__ double max_(double a, double b) { return fmax(a, b); }
  static __device__ __forceinline__ double fma_(double a, double b, double c) { return fma(a, b, c); }
  template <int N> static __device__ __forceinline__ double ipow(double x) {
    double r = x;
#pragma unroll
    for (int i = 1; i < N; ++i) r *= x;
    return r;
  }
  // |du| <= step_tol * min(u,1) ends the Halley iteration: the step AFTER that one is cubic in it
  static __device__ __forceinline__ double step_tol() { return 1e-6; }
  static __device__ __forceinline__ double nan_() { return __longlong_as_double(0x7ff8000000000000LL); }
  static __device__ __forceinline__ double inf_() { return __longlong_as_double(0x7ff0000000000000LL); }
};

template <> struct CatMath<float> {
  typedef float real;
  static __device__ __forceinline__ float sin_(float x) { return __sinf(x); }
  static __device__ __forceinline__ float cos_(float x) { return __cosf(x); }
  template <int N> static __device__ __forceinline__ void sin_v(float (&a)[N]) {
#pragma unroll
    for (int n = 0; n < N; ++n) a[n] = __sinf(a[n]);
  }
  static __device__ __forceinline__ float pos_(float x) { return x > 0.0f ? x : 0.0f; }
  static __device__ __forceinline__ void sincos_(float x, float* s, float* c) { __sincosf(x, s, c); }
  static __device__ __forceinline__ float tanh_(float x) {
    float r;
    asm("tanh.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
  }
  static __device__ __forceinline__ float exp_(float x) { return __expf(x); }
  static __device__ __forceinline__ float log_(float x) { return __logf(x); }
  // fast mode = single MUFU instructions (sqrt.approx / rcp.approx, <= 1-2 ulp), not the IEEE-rounded
  // sequences: __fsqrt_rn / __frcp_rn cost a MUFU + Newton step + a special-case branch each, which was 15% of
  // the fp32 rollout's instructions and most of its branches (profiles/r1_ncu_full_k_mpc_rollout_f32.csv)
  static __device__ __forceinline__ float sqrt_(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
  }
  static __device__ __forceinline__ float abs_(float x) { return fabsf(x); }
  static __device__ __forceinline__ float rcp_(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
  }
  static __device__ __forceinline__ float div_(float a, float b) { return __fdividef(a, b); }
  static __device__ __forceinline__ float rcp_fast_(float x) { return rcp_(x); }
  static __device__ __forceinline__ float min_(float a, float b) { return fminf(a, b); }
  static __device__ __forceinline__ float max_(float a, float b) { return fmaxf(a, b); }
  static __device__ __forceinline__ float clamp_sym_(float x, float c) { return fminf(fmaxf(x, -c), c); }
  static __device__ __forceinline__ float fma_(float a, float b, float c) { return fmaf(a, b, c); }
  template <int N> static __device__ __forceinline__ float ipow(float x) {
    float r = x;
#pragma unroll
    for (int i = 1; i < N; ++i) r *= x;
    return r;
  }
  static __device__ __forceinline__ float step_tol() { return 2e-3f; }
  static __device__ __forceinline__ float nan_() { return __int_as_float(0x7fc00000); }
  static __device__ __forceinline__ float inf_() { return __int_as_float(0x7f800000); }
};

// ------------------------------------------------------------------------------------------
// Catenary parameter.  Reference: main_fun.py:423  f(C) = C^2 (L^2 - dH^2) - 4 sinh^2(l C / 2).
// With u = C l / 2 and r = sqrt(L^2 - dH^2) / l the root satisfies sinh(u)/u = r.  We solve
//   g(u) = (sinh(u) - u) - e u = 0,   e = r - 1 > 0
// by Halley's method (g' = cosh(u) - 1 - e, g'' = sinh(u); one exp per iteration) and return
// sinh(u), cosh(u) at the root so the lowest point needs no further transcendental.
// ------------------------------------------------------------------------------------------
template <typename T> struct CatRoot {
  T u, sh, ch, cm;   // u = C l / 2, sinh(u), cosh(u), cosh(u) - 1
};

// warm start carried from the previous solve of a slowly moving span: the old root with
// sinh(u)-u, cosh(u)-1, sinh(u) there.  u == 0 means "none" (cold start).
template <typename T> struct CatWarm {
  T u, sm, cm, sh;
};

// sinh(u)-u, cosh(u)-1, sinh(u) for u > 0.
template <typename T, typename M>
__device__ __forceinline__ void cat_hyp(T u, T& sm, T& cm, T& sh) {
  if constexpr (sizeof(T) == 8) {
    cat_hyp_f64(u, sm, cm, sh);
  } else {
    if (u < T(1)) {   // series below 1 (no cancellation), exp above
      const T w = u * u;
      T p = M::fma_(w, T(1.0 / 110.0), T(1));
      p = M::fma_(w * T(1.0 / 72.0), p, T(1));
      p = M::fma_(w * T(1.0 / 42.0), p, T(1));
      p = M::fma_(w * T(1.0 / 20.0), p, T(1));
      sm = u * w * T(1.0 / 6.0) * p;
      T q = M::fma_(w, T(1.0 / 90.0), T(1));
      q = M::fma_(w * T(1.0 / 56.0), q, T(1));
      q = M::fma_(w * T(1.0 / 30.0), q, T(1));
      q = M::fma_(w * T(1.0 / 12.0), q, T(1));
      cm = w * T(0.5) * q;
      sh = sm + u;
    } else {
      const T E = M::exp_(u);
      const T Ei = M::rcp_(E);
      sh = T(0.5) * (E - Ei);
      sm = sh - u;
      cm = T(0.5) * (E + Ei) - T(1);
    }
  }
}

// cold-start guess, max relative error 6.4% over e in [1e-13, 1e13]
// A start value good to a few percent needs no fp64 library logarithm (3 x ~80 instructions per cold solve, and
// with e on both sides of 2 inside a warp both branches ran): single precision with the MUFU log/rcp/sqrt, both
// forms evaluated and selected (~25 instructions, no divergence).  The iteration converges to the same root.
template <typename T, typename M>
__device__ __forceinline__ T cat_u_guess(T e) {
  const float ef = fminf((float)e, 1e30f);   // e > 1e30 only for l < 1e-30, where C = 2u/l is outside [c_lo, c_hi]
  const float small = sqrtf(__fdividef(6.0f * ef, fmaf(0.3f, ef, 1.0f)));
  const float lr = __logf(2.0f * (1.0f + fmaxf(ef, 2.0f)));   // >= log 6 > 1
  const float large = lr + __logf(lr + __logf(lr));
  return T(ef < 2.0f ? small : large);
}

// (sinh u - u, cosh u - 1, sinh u) moved from u to u + d by the addition theorems, for |d| <= 0.03:
//   cosh d - 1 = z (1/2 + z (1/24 + z/720)),  sinh d - d = d z (1/6 + z (1/120 + z/5040)),  z = d^2
// (next terms z^4/40320 <= 1.7e-17 and d z^4/362880: below one ulp of the results).  17 flops against the 35
// plus exponent arithmetic of cat_hyp, no range reduction, and no cancellation: every product is added to a
// term of its own sign structure (sm absorbs sinh d - d directly).
__constant__ __align__(16) double CAT_SHIFT_C[6] = {1.0 / 720.0, 1.0 / 5040.0, 1.0 / 24.0, 1.0 / 120.0, 1.0 / 6.0, 0.0};
template <typename T> __device__ __forceinline__ T cat_shift_c(int i) {   // fp64: uniform loads, not literals
  if constexpr (sizeof(T) == 8) return CAT_SHIFT_C[i];
  else return i == 0 ? T(1.0 / 720.0) : i == 1 ? T(1.0 / 5040.0) : i == 2 ? T(1.0 / 24.0) : i == 3 ? T(1.0 / 120.0)
                                                                                                 : T(1.0 / 6.0);
}

template <typename T, typename M>
__device__ __forceinline__ void cat_hyp_shift(const CatWarm<T>& w, T d, T& sm, T& cm, T& sh) {
  const T z = d * d;
  const T cd = z * M::fma_(z, M::fma_(z, cat_shift_c<T>(0), cat_shift_c<T>(2)), T(0.5));              // cosh d - 1
  const T sd = d * z * M::fma_(z, M::fma_(z, cat_shift_c<T>(1), cat_shift_c<T>(3)), cat_shift_c<T>(4));   // sinh d - d
  const T shd = d + sd;
  const T ch = w.cm + T(1);
  sh = M::fma_(w.sh, cd, M::fma_(ch, shd, w.sh));
  cm = M::fma_(ch, cd, M::fma_(w.sh, shd, w.cm));
  sm = M::fma_(w.sh, cd, M::fma_(w.cm, shd, w.sm + sd));
}

// One safeguarded Halley step for the root of  G(u) = D (sinh u - u) - N u,  N = L^2 - dH^2 - l^2,
// D = l (sqrt(L^2 - dH^2) + l):  the same root as g(u) = (sinh u - u) - e u with e = N / D, without
// forming the quotient.  G' = D (cosh u - 1) - N, G'' = D sinh u.  The quotient of the step uses the
// fast reciprocal: an error of 1e-15 relative in a step that is itself a correction is invisible.
template <typename T, typename M, bool NEWTON = false>
__device__ __forceinline__ T cat_halley_step(T u, T N, T D, T sm, T cm, T sh) {
  const T G = D * sm - N * u;
  const T Gp = D * cm - N;
  // Branch-free: the cases share one reciprocal (selected operands, identical arithmetic per lane), so the
  // step is straight-line code inside the caller's basic block.
  const bool newton = Gp > T(0);
  T du;
  if constexpr (NEWTON) {
    // polishing step next to the root (the rollout after its warm Halley step): Newton's quadratic
    // convergence from ~1e-9 is already below one ulp, at a third of the arithmetic
    du = G * M::rcp_fast_(Gp);
  } else {
    const T den = T(2) * Gp * Gp - G * (D * sh);
    const bool halley = newton && den > T(0);
    const T num = halley ? T(2) * G * Gp : G;
    const T dn = halley ? den : Gp;
    du = num * M::rcp_fast_(dn);               // Halley, else Newton (junk when !newton, replaced below)
  }
  if (!newton) du = M::fma_(T(-0.5), u, T(-0.1));   // left of the minimum of G: move right
  if (!(du < u)) du = T(0.5) * u;              // keep u > 0
  return du;
}

// N > 0, D > 0 required.  With a warm start the first Halley step costs no transcendental (the carried
// sinh/cosh are exact at the old root).  A rollout step moves the span by ~1e-3 relative, so that
// step lands within ~1e-9 of the root and ONE full iteration finishes; the loop still runs until
// |du| is below tol.  Returns the root with sinh-u, cosh-1, sinh Taylor-corrected to it.
template <typename T> struct CatSolve {
  T u, sh, ch;       // the root, sinh and cosh there
  CatWarm<T> rec;    // record to hand to the next solve of the same span
};

// the iteration loop of the rollout's solve (cold starts, fast-moving spans), out of line: it keeps the loop body of
// the rollout small, and its cost does not matter where it runs
template <typename T> struct CatIter {
  T u, sm, cm, sh, du;
};
template <typename T, typename M>
__device__ __noinline__ CatIter<T> cat_solve_iterate(T u, T N, T D, bool more) {
  CatIter<T> r;
  r.u = u;
  r.sm = r.cm = r.sh = r.du = T(0);
  int it = 1;
  // at least one evaluation: the caller's (sm, cm, sh) belong to the point before its last step
  do {
    cat_hyp<T, M>(r.u, r.sm, r.cm, r.sh);
    r.du = cat_halley_step<T, M>(r.u, N, D, r.sm, r.cm, r.sh);
    r.u -= r.du;
    more = M::abs_(r.du) > M::step_tol() * (r.u < T(1) ? r.u : T(1));
    ++it;
  } while (more && it < 60);
  return r;
}

template <typename T, typename M, bool FULL_WARP = false>
__device__ __forceinline__ CatSolve<T> cat_solve_u(T N, T D, const CatWarm<T>& w) {
  T u;
  T sm, cm, sh, du;
  bool far = false;
  // The first iteration is straight-line code: with a warm start it is also the last one, and without a
  // data-dependent loop around it the compiler can interleave this chain with the caller's independent work
  // (the RK4 stages of the rollout).  Further iterations (cold starts, fast-moving spans) take the loop.
  if constexpr (FULL_WARP) {   // callers always hand in a record (cat_warm_guess before their loop)
    // the warm step moves u by d ~ 1e-3 u: the record follows by the addition theorems instead of a fresh
    // sinh/cosh evaluation; a step beyond their range is sent to the loop, which starts with cat_hyp(u)
    const T d = -cat_halley_step<T, M>(w.u, N, D, w.sm, w.cm, w.sh);
    u = w.u + d;
    cat_hyp_shift<T, M>(w, d, sm, cm, sh);
    far = !(M::abs_(d) <= T(0.03));
    du = cat_halley_step<T, M, true>(u, N, D, sm, cm, sh);
  } else {
    if (w.u > T(0)) u = w.u - cat_halley_step<T, M>(w.u, N, D, w.sm, w.cm, w.sh);
    else u = cat_u_guess<T, M>(M::div_(N, D));
    cat_hyp<T, M>(u, sm, cm, sh);
    du = cat_halley_step<T, M>(u, N, D, sm, cm, sh);
  }
  u -= du;
  // |du| > tol * min(u, 1)  <=>  |du| > tol * u  or  |du| > tol
  bool more = M::abs_(du) > M::step_tol() * u || M::abs_(du) > M::step_tol() || far;
  bool again;
  if constexpr (FULL_WARP) again = __any_sync(0xffffffffu, more);   // every lane of the warp is here
  else again = more;
  if (again) {
    if constexpr (FULL_WARP) {
      cat_dbg_count(0);
      if (more) {   // only the lanes that need it iterate (out of line); the others keep their evaluation
        const CatIter<T> r = cat_solve_iterate<T, M>(u, N, D, more);
        u = r.u;
        sm = r.sm;
        cm = r.cm;
        sh = r.sh;
        du = r.du;
      }
    } else {
      int it = 1;
      while (more && it < 60) {
        cat_hyp<T, M>(u, sm, cm, sh);
        du = cat_halley_step<T, M>(u, N, D, sm, cm, sh);
        u -= du;
        more = M::abs_(du) > M::step_tol() * (u < T(1) ? u : T(1));
        ++it;
      }
    }
  }
  // (sm, cm, sh) are values at u + du, the point of the last evaluation; |du| <= step_tol * min(u, 1).
  // d/du: sm' = cm, cm' = sh, sh' = ch
  const T ch = cm + T(1);
  CatSolve<T> r;
  r.u = u;
  if constexpr (FULL_WARP) {
    // outputs to first order (remainder du^2/2 <= 5e-13 relative); the record stays at the evaluation point,
    // one polishing step (~1e-9 u) behind the root -- as good a start for the next warm step as the root
    r.sh = M::fma_(-ch, du, sh);
    r.ch = M::fma_(-sh, du, ch);
    r.rec.u = u + du;
    r.rec.sm = sm;
    r.rec.cm = cm;
    r.rec.sh = sh;
  } else {
    // second order (remainder du^3/6 < 2e-19): the record is moved to the root
    const T h2 = du * du * T(0.5);
    r.rec.u = u;
    r.rec.sm = M::fma_(h2, sh, M::fma_(-cm, du, sm));
    r.rec.cm = M::fma_(h2, ch, M::fma_(-sh, du, cm));
    r.rec.sh = M::fma_(h2, sh, M::fma_(-ch, du, sh));
    r.sh = r.rec.sh;
    r.ch = r.rec.cm + T(1);
  }
  return r;
}

// Catenary terms of one span: horizontal l, vertical dH = zB - zA, cable L.
// C is NaN outside the reference's domain: brentq on [c_lo, c_hi] (main_fun.py:425) raises when
// f(c_lo), f(c_hi) have the same sign -> nan (main_fun.py:427-428).  f > 0 exactly on (0, root),
// so the sign test is c_lo < root < c_hi with a root existing iff L^2 - dH^2 > l^2.
// z_low_rel is the lowest point relative to the mean end height:
//   z_low = (zA+zB)/2 - (L/2) coth(u) + l/(2u)   if |dH| <= L tanh(u)  (vertex inside the span,
//   equivalent to 0 <= l/2 - atanh(dH/L)/C <= l), else min(zA, zB).
template <typename T> struct CatSpan {
  T C;        // NaN when outside the domain (NEED_C false: 0 inside the domain, NaN outside)
  T u;        // root (valid even if C is out of [c_lo,c_hi]); 0 when no root exists
  T sh, ch;
  bool inside;
  CatWarm<T> warm;   // hand to the next solve of the same (slowly moving) span
};

template <typename T> __device__ __forceinline__ CatWarm<T> cat_cold() {
  CatWarm<T> w;
  w.u = T(0);
  w.sm = T(0);
  w.cm = T(0);
  w.sh = T(0);
  return w;
}

// What the solve needs from a span: s2 = L^2 - dH^2, N = s2 - l^2, D = l (sqrt(s2) + l).
template <typename T> struct CatSpanPrep {
  T N, D;
  bool none;   // no root: taut, vertical or degenerate span
};

template <typename T, typename M>
__device__ __forceinline__ CatSpanPrep<T> cat_span_prep(T l, T dH, T L) {
  CatSpanPrep<T> p;
  const T s2 = (L - dH) * (L + dH);
  const T sq = M::sqrt_(M::pos_(s2));
  p.N = s2 - l * l;
  p.D = l * (sq + l);
  p.none = !(s2 > T(0)) || !(p.N > T(0)) || !(l > T(0));
  return p;
}

// fp32 fast mode, stand-alone solves and shapes: N = L^2 - dH^2 - l^2 is formed in fp64 from the fp32 inputs
// (l2 = l^2 or rel_x^2 + rel_y^2, exact in fp64) and rounded ONCE.  In fp32 the two subtractions lose
// eps * L^2 / N relative -- 3e-4 in C for a span within 0.1% of taut -- while everything after N (the quotient-free
// iteration, D, the outputs) is benign in 24 bits.  Four FP64 instructions per solve.
template <typename M>
__device__ __forceinline__ CatSpanPrep<float> cat_span_prep_exact(float l, double l2, double dH, double L) {
  CatSpanPrep<float> p;
  const double s2d = (L - dH) * (L + dH);
  const float s2 = (float)s2d;
  const float sq = M::sqrt_(M::pos_(s2));
  p.N = (float)(s2d - l2);
  p.D = l * (sq + l);
  p.none = !(s2d > 0.0) || !(s2d - l2 > 0.0) || !(l > 0.0f);
  return p;
}

// NEED_C = false skips the division C = 2u/l: the domain test is done on 2u against c*l.
template <typename T, typename M, bool NEED_C = true, bool FULL_WARP = false>
__device__ __forceinline__ CatSpan<T> cat_span_from(const CatSpanPrep<T>& p, T l, T dH, T L, T c_lo, T c_hi,
                                                    const CatWarm<T>& warm) {
  CatSpan<T> s;
  const bool none = p.none;
  if constexpr (!FULL_WARP) {
    if (none) {
      s.C = M::nan_();
      s.u = T(0);
      s.sh = T(0);
      s.ch = T(1);
      s.inside = false;
      s.warm = cat_cold<T>();
      return s;
    }
  }
  // FULL_WARP (the rollout): no early exit -- a rootless lane solves a harmless stand-in (N = D = 1) and is
  // masked afterwards, so the warp stays converged and the common path is one basic block
  const CatSolve<T> r =
      cat_solve_u<T, M, FULL_WARP>(FULL_WARP && none ? T(1) : p.N, FULL_WARP && none ? T(1) : p.D, warm);
  s.u = r.u;
  s.sh = r.sh;
  s.ch = r.ch;
  s.warm = r.rec;
  if constexpr (NEED_C) {
    const T C = M::div_(T(2) * r.u, l);
    s.C = (C > c_lo && C < c_hi) ? C : M::nan_();
  } else {
    const T tu = T(2) * r.u;
    s.C = (tu > c_lo * l && tu < c_hi * l) ? T(0) : M::nan_();
  }
  s.inside = M::abs_(dH) * s.ch <= L * r.sh;
  // FULL_WARP: only C is masked.  u, sh, ch of a rootless lane are the stand-in's and are never read behind a
  // NaN C (cat_z_low, the penalty); its warm record stays the stand-in root -- a consistent (u, sinh, cosh)
  // triple from which the safeguarded iteration converges like from any other start.
  if constexpr (FULL_WARP) s.C = none ? M::nan_() : s.C;
  return s;
}

template <typename T, typename M, bool NEED_C = true, bool FULL_WARP = false>
__device__ __forceinline__ CatSpan<T> cat_span(T l, T dH, T L, T c_lo, T c_hi, const CatWarm<T>& warm) {
  return cat_span_from<T, M, NEED_C, FULL_WARP>(cat_span_prep<T, M>(l, dH, L), l, dH, L, c_lo, c_hi, warm);
}

// a span given by (l, dH) as main_fun.solve_catenary takes it
template <typename T, typename M>
__device__ __forceinline__ CatSpan<T> cat_span_ldh(T l, T dH, T L, T c_lo, T c_hi) {
  if constexpr (sizeof(T) == 4) {
    return cat_span_from<T, M>(cat_span_prep_exact<M>(l, (double)l * (double)l, (double)dH, (double)L), l, dH, L,
                               c_lo, c_hi, cat_cold<T>());
  } else {
    return cat_span<T, M>(l, dH, L, c_lo, c_hi, cat_cold<T>());
  }
}

// Warm-start record at the cold-start guess of a span, without solving: (u, sinh u - u, cosh u - 1, sinh u)
// consistent at u.  The rollout builds it once per candidate from the initial pose, so every solve inside its
// loop is the warm, branch-free form.
template <typename T, typename M>
__device__ __forceinline__ CatWarm<T> cat_warm_guess(T l, T dH, T L) {
  const T s2 = (L - dH) * (L + dH);
  const T sq = M::sqrt_(M::pos_(s2));
  T N = s2 - l * l;
  T D = l * (sq + l);
  if (!(s2 > T(0)) || !(N > T(0)) || !(l > T(0))) {
    N = T(1);
    D = T(1);
  }
  CatWarm<T> w;
  w.u = cat_u_guess<T, M>(M::div_(N, D));
  cat_hyp<T, M>(w.u, w.sm, w.cm, w.sh);
  return w;
}

// lowest z of the span (world frame), SURVEY.md 8a "per-step catenary terms"
template <typename T, typename M>
__device__ __forceinline__ T cat_z_low(const CatSpan<T>& s, T l, T L, T zA, T zB) {
  const T zmin = zB < zA ? zB : zA;
  // (zA+zB)/2 + (l sh - L ch u) / (2 sh u); select instead of an early return (no branch in the rollout loop)
  const T zv = T(0.5) * (zA + zB) + (l * s.sh - L * s.ch * s.u) * M::rcp_fast_(T(2) * s.sh * s.u);
  return (!(s.C == s.C) || !s.inside) ? zmin : zv;
}

// ------------------------------------------------------------------------------------------
// Rodrigues rotation, main_fun.py:18-35 (axis re-normalised on every call, as the reference does)
// ------------------------------------------------------------------------------------------
template <typename T, typename M>
__device__ __forceinline__ void cat_rodrigues(const T v[3], const T axis_in[3], T c, T s, T out[3]) {
  const T n = M::sqrt_(axis_in[0] * axis_in[0] + axis_in[1] * axis_in[1] + axis_in[2] * axis_in[2]);
  const T k[3] = {M::div_(axis_in[0], n), M::div_(axis_in[1], n), M::div_(axis_in[2], n)};
  const T cx = k[1] * v[2] - k[2] * v[1];
  const T cy = k[2] * v[0] - k[0] * v[2];
  const T cz = k[0] * v[1] - k[1] * v[0];
  const T dot = k[0] * v[0] + k[1] * v[1] + k[2] * v[2];
  const T omc = T(1) - c;
  out[0] = v[0] * c + cx * s + k[0] * dot * omc;
  out[1] = v[1] * c + cy * s + k[1] * dot * omc;
  out[2] = v[2] * c + cz * s + k[2] * dot * omc;
}

// theta rotation axis, main_fun.py:75-89 (fallbacks below 1e-9)
template <typename T, typename M>
__device__ __forceinline__ void cat_theta_axis(const T conn[3], T axis[3]) {
  T px = conn[0], py = conn[1];
  const T n = M::sqrt_(px * px + py * py);
  if (n < T(1e-9)) {
    px = T(1);
    py = T(0);
  } else {
    px = M::div_(px, n);
    py = M::div_(py, n);
  }
  // cross((px,py,0),(0,0,1)) = (py, -px, 0)
  T ax = py, ay = -px;
  const T n2 = M::sqrt_(ax * ax + ay * ay);
  if (n2 < T(1e-9)) {
    axis[0] = T(0);
    axis[1] = T(1);
    axis[2] = T(0);
  } else {
    axis[0] = M::div_(ax, n2);
    axis[1] = M::div_(ay, n2);
    axis[2] = T(0);
  }
}

// ------------------------------------------------------------------------------------------
// 18-column features, simple.py:26-41, then StandardScaler (test_temp.py:49).
// All 18 are written; the generated equations read a few and the compiler drops the rest.
// ------------------------------------------------------------------------------------------
template <typename T, typename M>
__device__ __forceinline__ void cat_features18(const T P0[3], const T P1[3], const T V1[3], const T A1[3],
                                               T th, T ga, T thp, T gap, T (&x)[18]) {
  const T rx = P1[0] - P0[0], ry = P1[1] - P0[1], rz = P1[2] - P0[2];
  const T nrm = M::sqrt_(rx * rx + ry * ry + rz * rz);
  const T inv = M::rcp_(nrm + T(1e-8));
  const T ux = rx * inv, uy = ry * inv, uz = rz * inv;
  const T nv = M::sqrt_(V1[0] * V1[0] + V1[1] * V1[1] + V1[2] * V1[2]) + T(1e-8);
  const T dot = V1[0] * ux + V1[1] * uy + V1[2] * uz;
  x[0] = P1[0]; x[1] = P1[1]; x[2] = P1[2];
  x[3] = V1[0]; x[4] = V1[1]; x[5] = V1[2];
  x[6] = A1[0]; x[7] = A1[1]; x[8] = A1[2];
  x[9] = ux; x[10] = uy; x[11] = uz;
  x[12] = M::min_(M::max_(nrm, T(1e-5)), T(10));
  x[13] = M::min_(M::max_(M::div_(dot, nv), T(-1)), T(1));
  x[14] = th; x[15] = ga; x[16] = thp; x[17] = gap;
}

// ------------------------------------------------------------------------------------------
// (cost, index) pairs with np.argmin ordering: NaN beats everything, then lower cost, then
// lower index (first minimum; fully_augmented_catenary.py:21 semantics).
// ------------------------------------------------------------------------------------------
struct CatBest {
  double cost;
  long long idx;
};

__host__ __device__ __forceinline__ bool cat_better(double ca, long long ia, double cb, long long ib) {
  const bool na = ca != ca, nb = cb != cb;
  if (na != nb) return na;
  if (na) return ia < ib;
  if (ca < cb) return true;
  if (ca > cb) return false;
  return ia < ib;
}

// Winner of a sampling-MPC step: lower cost, then lower index, and a NaN cost ranks with +inf -- a candidate
// whose rollout produced a NaN must never become the applied control (a receding-horizon loop would feed it to
// the plant).  With every cost NaN the lowest index wins, which in the controller is candidate 0, the noise-free
// shifted previous solution.  Definition: oracle.argmin_mpc.  (np.argmin's NaN-first order above stays for
// the curve's lowest sample, fully_augmented_catenary.py:21.)
__host__ __device__ __forceinline__ bool cat_better_mpc(double ca, long long ia, double cb, long long ib) {
  const double inf = __builtin_huge_val();
  const double a = (ca != ca) ? inf : ca, b = (cb != cb) ? inf : cb;
  if (a < b) return true;
  if (a > b) return false;
  return ia < ib;
}

// ------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011) -- counter-based RNG for on-device control sampling
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void cat_philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                  uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ k1;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// ------------------------------------------------------------------------------------------
// Standard normal from one Philox word by a tabulated inverse CDF (gen/icdf_table.cuh; the table is the
// definition, shared with oracle.sample_controls -- tools/make_icdf_table.py):
//   z(r) = sign * y(r'),  r' = r < 2^31 ? r : ~r (clamped to >= 32),  y(r') = -ndtri((r' + 0.5) / 2^32)
// y is a cubic in the bits below the top six of r' -- octave e = floor(log2 r') and the next five bits select one
// of 32 segments per octave, exactly the exponent and leading mantissa bits of (double)r'.  Per normal: one
// conversion, five integer ops, two 16-byte shared loads, three FMAs -- a quarter of a Box-Muller lane with its
// logarithm, square root and sine/cosine (the round-1 sampler: ~230 of the 642 instructions per warp-step).
// Shared memory holds octaves >= 21 (|z| <= 3.297, 10 KB in fp64): candidates are clipped at +-clip (3 sigma
// in every BASELINE config), so below that octave the lane is clamped and needs no coefficients; a caller that
// asks for clip > 3.29 sends those lanes (2^-10 of them) to the full table in global memory.
// ------------------------------------------------------------------------------------------
#include "gen/icdf_table.cuh"
#define CAT_ICDF_SMEM_E_LO 21
#define CAT_ICDF_SMEM_SEGS ((CAT_ICDF_E_HI - CAT_ICDF_SMEM_E_LO + 1) * 32)
#define CAT_ICDF_SMEM_ZMAX 3.29   // < y(2^21) = 3.2971...

template <typename T> struct CatIcdfSmem {
  alignas(16) T tab[(CAT_ICDF_SMEM_SEGS + 1) * 4];   // row 0 = (4, 0, 0, 0): every octave below the hot ones
};

// block-cooperative copy of the hot octaves into shared memory; ends with a barrier
template <typename T>
__device__ __forceinline__ void cat_icdf_stage(CatIcdfSmem<T>& sm) {
  const double2* src =
      reinterpret_cast<const double2*>(CAT_ICDF_TAB + (CAT_ICDF_SMEM_E_LO - CAT_ICDF_E_LO) * 32 * 4);
  for (int i = threadIdx.x; i < CAT_ICDF_SMEM_SEGS * 2; i += blockDim.x) {
    const double2 v = __ldg(src + i);
    sm.tab[4 + 2 * i] = T(v.x);
    sm.tab[4 + 2 * i + 1] = T(v.y);
  }
  if (threadIdx.x < 4) sm.tab[threadIdx.x] = threadIdx.x == 0 ? T(4) : T(0);   // > every clip <= CAT_ICDF_SMEM_ZMAX
  __syncthreads();
}

// The table is addressed through its 32-bit shared-window offset, taken ONCE per kernel: with a generic pointer the
// compiler re-derived the window base (S2UR + UMOV + ULEA) in front of every lookup -- three issue slots per normal.
typedef unsigned CatIcdfRef;
template <typename T> __device__ __forceinline__ CatIcdfRef cat_icdf_ref_of(const CatIcdfSmem<T>& sm) {
  return (unsigned)__cvta_generic_to_shared(sm.tab);
}

// lanes below the shared-memory octaves when the caller's clip reaches past them: the full table in global memory.
// Kept out of line so that the common path carries one predicated call, not the predicated-off loads and FMAs
// (round-2 ncu: ~32 of the sampler's instructions per warp-step were this never-taken block, if-converted).
template <typename T, typename M>
__device__ __noinline__ T cat_icdf_deep(int seg, T t) {
  const double* g = CAT_ICDF_TAB + 4 * seg;
  return M::fma_(M::fma_(M::fma_(T(__ldg(g + 3)), t, T(__ldg(g + 2))), t, T(__ldg(g + 1))), t, T(__ldg(g)));
}

// clip(z(r), +-clip);  DEEP: clip > CAT_ICDF_SMEM_ZMAX, lanes below the hot octaves read the full table
template <typename T, typename M, bool DEEP>
__device__ __forceinline__ T cat_icdf_normal(uint32_t r, CatIcdfRef tab, T clip, unsigned one_hi = 0x3ff00000u) {
  const uint32_t flip = (uint32_t)((int)r >> 31);          // 0xffffffff for r >= 2^31
  uint32_t rp = r ^ flip;                                   // r' = r < 2^31 ? r : ~r
  // the clamp to 32 matters where the segment indexes the full table; on the shared table every r' below the hot
  // octaves (0 included: its exponent field is 0) lands on row 0 through the row clamp further down
  if constexpr (DEEP) rp = rp < 32u ? 32u : rp;
  int seg;
  T t;
  if constexpr (sizeof(T) == 8) {
    const double x = (double)rp;   // exact (rp < 2^31)
    const int hi = __double2hiint(x);
    const unsigned lo = (unsigned)__double2loint(x);
    seg = (hi >> 15) - ((1023 + CAT_ICDF_E_LO) << 5);
    // mantissa bits below the top five, as 1.t - 1
    unsigned mh;   // (funnel & 0x000fffff) | 0x3ff00000 as ONE three-input logic op (the compiler emits two)
    asm("lop3.b32 %0, %1, 0x000fffff, %2, 0xea;" : "=r"(mh) : "r"(__funnelshift_l(lo, (unsigned)hi, 5)), "r"(one_hi));
    t = __hiloint2double((int)mh, (int)(lo << 5)) - 1.0;
  } else {
    int b = __float_as_int((float)rp);   // 24-bit rounding moves the sample by < 1e-6 ...
    b = b > 0x4effffff ? 0x4effffff : b;   // ... and must not round up to 2^31, one octave past the table
    seg = (b >> 18) - ((127 + CAT_ICDF_E_LO) << 5);
    t = __int_as_float(0x3f800000 | ((b & 0x3ffff) << 5)) - 1.0f;
  }
  // row of the shared table: 1 + segment above the first hot octave; everything below lands on row 0, whose constant
  // 4 exceeds every clip the shared table serves, so the clamp below turns it into +-clip without a select here
  const int row = seg - ((CAT_ICDF_SMEM_E_LO - CAT_ICDF_E_LO) * 32 - 1);
  const unsigned at = tab + (unsigned)(row < 0 ? 0 : row) * (unsigned)(4 * sizeof(T));
  T c0, c1, c2, c3;   // volatile: must stay behind the barrier that ends the table's staging
  if constexpr (sizeof(T) == 8) {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(c0), "=d"(c1) : "r"(at));
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+16];" : "=d"(c2), "=d"(c3) : "r"(at));
  } else {
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(c0), "=f"(c1), "=f"(c2), "=f"(c3) : "r"(at));
  }
  T y = M::fma_(M::fma_(M::fma_(c3, t, c2), t, c1), t, c0);
  if constexpr (DEEP) {
    if (row <= 0) y = cat_icdf_deep<T, M>(seg, t);
  }
  // |z| clipped BEFORE the sign goes on (y >= 0, never NaN: one compare and two selects), then z = r >= 2^31 ? y : -y
  // on the sign bit
  y = y > clip ? clip : y;
  if constexpr (sizeof(T) == 8)
    return __hiloint2double(__double2hiint(y) ^ (int)(~flip & 0x80000000u), __double2loint(y));
  else
    return __int_as_float(__float_as_int(y) ^ (int)(~flip & 0x80000000u));
}

template <typename T, typename M>
__device__ __noinline__ T cat_icdf_normal_deep(uint32_t r, CatIcdfRef tab, T clip) {   // out of line: rare, and small
  return cat_icdf_normal<T, M, true>(r, tab, clip);
}

// three clipped normals scaled by sigma, keyed by (seed, problem, global k, h): Philox words 0..2 -> axes x, y, z.
// The deep-tail form (clip beyond the shared octaves; no BASELINE config) is a separate instance behind ONE
// warp-uniform branch per call, so the common one carries no per-normal test, branch or reconvergence point.
template <typename T, typename M>
__device__ __forceinline__ void cat_sample_control(uint64_t seed, uint32_t problem, uint64_t k, uint32_t h,
                                                   const T sigma[3], T clip, bool deep, CatIcdfRef sm, T out[3]) {
  uint32_t r[4];
  cat_philox4x32_10((uint32_t)k, (uint32_t)(k >> 32), h, problem, (uint32_t)seed, (uint32_t)(seed >> 32), r);
  if (deep) {   // clip > CAT_ICDF_SMEM_ZMAX, decided by the host: a uniform integer test the compiler keeps out of the loop
#pragma unroll
    for (int a = 0; a < 3; ++a) out[a] = cat_icdf_normal_deep<T, M>(r[a], sm, clip) * sigma[a];
  } else {
#pragma unroll
    for (int a = 0; a < 3; ++a) out[a] = cat_icdf_normal<T, M, false>(r[a], sm, clip) * sigma[a];
  }
}

// A loop-invariant the compiler must not re-derive inside the loop: the value goes through an opaque move, so it
// occupies registers instead of being rebuilt from the constant bank / special registers at every use (the sampler's
// table address cost S2UR + UMOV + ULEA per lookup, its clip two LDCU per normal).
__device__ __forceinline__ unsigned cat_pin(unsigned v) {
  unsigned o;
  asm volatile("mov.u32 %0, %1;" : "=r"(o) : "r"(v));
  return o;
}
__device__ __forceinline__ double cat_pin(double v) {
  double o;
  asm volatile("mov.f64 %0, %1;" : "=d"(o) : "d"(v));
  return o;
}
__device__ __forceinline__ float cat_pin(float v) {
  float o;
  asm volatile("mov.f32 %0, %1;" : "=f"(o) : "f"(v));
  return o;
}
