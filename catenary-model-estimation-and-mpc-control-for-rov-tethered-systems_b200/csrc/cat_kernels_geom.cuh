// cat_kernels_geom.cuh -- batched catenary parameter solve and curve-shape kernels (sm_100a).
#pragma once
#include "cat_math.cuh"

// ------------------------------------------------------------------------------------------
// solve_catenary(l, delta_H, L)  -- main_fun.py:418-431.  One thread per configuration,
// SoA inputs, coalesced.  8 B (fp64) in x2(+1), 8 B out per config: compute bound.
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_solve_param(long long n, const T* __restrict__ l,
                                                     const T* __restrict__ dH, const T* __restrict__ L,
                                                     long long L_stride, T c_lo, T c_hi,
                                                     T* __restrict__ C_out) {
  typedef CatMath<T> M;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const CatSpan<T> s = cat_span<T, M>(l[i], dH[i], L[i * L_stride], c_lo, c_hi, cat_cold<T>());
  C_out[i] = s.C;
}

// ------------------------------------------------------------------------------------------
// Curve shapes: Catenary.__call__ (pympc; builder-defined sampling, see oracle.catenary_points)
// fused with transform_catenary (main_fun.py:38-111).
//
// Block = CAT_SHAPE_TILE configurations.  Phase 1: thread t solves configuration t of the tile
// (both spans A->B and A->B') and parks the per-span constants in shared memory.  Phase 2: the
// block walks the tile's n_pts*TILE points with the point index fastest, so a warp writes 32
// consecutive (x,y,z) triples = 768 contiguous bytes of the AoS (n, n_pts, 3) output.
// ------------------------------------------------------------------------------------------
#define CAT_SHAPE_TILE 128

template <typename T> struct CatShapeSpan {
  T rel[3];     // B - A
  T l;          // horizontal span
  T two_u;      // C * l
  T v_low;      // C * x_low = u - atanh(dH/L)
  T inv_C;      // 1/C
  T ch_low;     // cosh(v_low)
  T Bend[3];    // exact end point (chord fallback pins it)
  int valid;
};

template <typename T> struct CatShapeCfg {
  T A[3];
  CatShapeSpan<T> sp[2];
  T k_theta[3], k_gamma[3];
  T ct, st, cg, sg;
};

template <typename T, typename M>
__device__ __forceinline__ void cat_shape_span_setup(const T A[3], const T Bq[3], T L, T c_lo, T c_hi,
                                                     CatShapeSpan<T>& sp, T* C_out) {
  sp.rel[0] = Bq[0] - A[0];
  sp.rel[1] = Bq[1] - A[1];
  sp.rel[2] = Bq[2] - A[2];
  sp.Bend[0] = Bq[0];
  sp.Bend[1] = Bq[1];
  sp.Bend[2] = Bq[2];
  const T l = M::sqrt_(sp.rel[0] * sp.rel[0] + sp.rel[1] * sp.rel[1]);
  const T dH = sp.rel[2];
  const CatSpan<T> s = cat_span<T, M>(l, dH, L, c_lo, c_hi, cat_cold<T>());
  sp.l = l;
  sp.valid = (s.C == s.C) ? 1 : 0;
  if (C_out) *C_out = s.C;
  if (sp.valid) {
    sp.two_u = T(2) * s.u;
    sp.v_low = s.u - T(0.5) * M::log_(M::div_(L + dH, L - dH));
    sp.inv_C = M::rcp_(s.C);
    const T E = M::exp_(sp.v_low);
    sp.ch_low = T(0.5) * (E + M::rcp_(E));
  } else {
    sp.two_u = T(0);
    sp.v_low = T(0);
    sp.inv_C = T(0);
    sp.ch_low = T(1);
  }
}

template <typename T, typename M>
__device__ __forceinline__ void cat_shape_point(const T A[3], const CatShapeSpan<T>& sp, T s, bool last, T p[3]) {
  if (s == T(0)) {          // the curve starts exactly at A (cosh is even: the oracle's z_0 - zA is exactly 0)
    p[0] = A[0];
    p[1] = A[1];
    p[2] = A[2];
  } else if (sp.valid) {
    p[0] = A[0] + sp.rel[0] * s;
    p[1] = A[1] + sp.rel[1] * s;
    const T t = sp.two_u * s - sp.v_low;
    const T E = M::exp_(t);
    const T ch = T(0.5) * (E + M::rcp_(E));
    p[2] = A[2] + (ch - sp.ch_low) * sp.inv_C;
  } else if (last) {
    p[0] = sp.Bend[0];
    p[1] = sp.Bend[1];
    p[2] = sp.Bend[2];
  } else {
    p[0] = A[0] + sp.rel[0] * s;
    p[1] = A[1] + sp.rel[1] * s;
    p[2] = A[2] + sp.rel[2] * s;
  }
}

template <typename T>
__global__ void __launch_bounds__(CAT_SHAPE_TILE) k_shape(long long n, const T* __restrict__ A,
                                                          const T* __restrict__ B, const T* __restrict__ L,
                                                          long long L_stride, const T* __restrict__ theta,
                                                          const T* __restrict__ gamma, int n_pts, int which,
                                                          T c_lo, T c_hi, T* __restrict__ out,
                                                          T* __restrict__ C_out, unsigned char* __restrict__ valid) {
  typedef CatMath<T> M;
  __shared__ CatShapeCfg<T> cfg[CAT_SHAPE_TILE];
  const long long base = (long long)blockIdx.x * CAT_SHAPE_TILE;
  const int tid = threadIdx.x;
  const long long i = base + tid;
  const bool need_rot = (which & (CAT_OUT_THETA_ROTATED | CAT_OUT_THETA_ALIGNED | CAT_OUT_FINAL)) != 0;
  if (i < n) {
    CatShapeCfg<T>& c = cfg[tid];
    const T a[3] = {A[3 * i], A[3 * i + 1], A[3 * i + 2]};
    const T b[3] = {B[3 * i], B[3 * i + 1], B[3 * i + 2]};
    const T Li = L[i * L_stride];
    c.A[0] = a[0];
    c.A[1] = a[1];
    c.A[2] = a[2];
    T Cv;
    cat_shape_span_setup<T, M>(a, b, Li, c_lo, c_hi, c.sp[0], &Cv);
    if (C_out) C_out[i] = Cv;
    int vbits = c.sp[0].valid;
    if (need_rot) {
      const T th = theta ? theta[i] : T(0);
      const T ga = gamma ? gamma[i] : T(0);
      M::sincos_(th, &c.st, &c.ct);
      M::sincos_(ga, &c.sg, &c.cg);
      const T conn[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]};
      T kt[3];
      cat_theta_axis<T, M>(conn, kt);
      // main_fun.py:92  rotated_B = A + rodrigues(conn, k_theta, theta)
      T rb[3];
      cat_rodrigues<T, M>(conn, kt, c.ct, c.st, rb);
      const T bq[3] = {a[0] + rb[0], a[1] + rb[1], a[2] + rb[2]};
      cat_shape_span_setup<T, M>(a, bq, Li, c_lo, c_hi, c.sp[1], (T*)nullptr);
      vbits |= c.sp[1].valid << 1;
      // axes as rodrigues_rotation normalises them (main_fun.py:30)
      const T nt = M::sqrt_(kt[0] * kt[0] + kt[1] * kt[1] + kt[2] * kt[2]);
      c.k_theta[0] = M::div_(kt[0], nt);
      c.k_theta[1] = M::div_(kt[1], nt);
      c.k_theta[2] = M::div_(kt[2], nt);
      // main_fun.py:102-103 gamma axis = conn/|conn| (then normalised again inside rodrigues)
      const T nc = M::sqrt_(conn[0] * conn[0] + conn[1] * conn[1] + conn[2] * conn[2]);
      const T kg[3] = {M::div_(conn[0], nc), M::div_(conn[1], nc), M::div_(conn[2], nc)};
      const T ng = M::sqrt_(kg[0] * kg[0] + kg[1] * kg[1] + kg[2] * kg[2]);
      c.k_gamma[0] = M::div_(kg[0], ng);
      c.k_gamma[1] = M::div_(kg[1], ng);
      c.k_gamma[2] = M::div_(kg[2], ng);
    }
    if (valid) valid[i] = (unsigned char)vbits;
  }
  __syncthreads();

  const long long n_here = (n - base < CAT_SHAPE_TILE) ? (n - base) : CAT_SHAPE_TILE;
  const long long total = n_here * n_pts;
  const T inv_nm1 = (n_pts > 1) ? M::rcp_(T(n_pts - 1)) : T(0);
  const long long blk = n * (long long)n_pts * 3;   // elements per output block
  for (long long j = tid; j < total; j += CAT_SHAPE_TILE) {
    const int ci = (int)(j / n_pts);
    const int pt = (int)(j - (long long)ci * n_pts);
    const CatShapeCfg<T>& c = cfg[ci];
    // s = pt/(n-1) as the oracle computes it (division), end point exactly 1
    const T s = (pt == n_pts - 1) ? T(1) : M::div_(T(pt), T(n_pts - 1));
    (void)inv_nm1;
    const bool last = pt == n_pts - 1;
    T* o = out + ((base + ci) * (long long)n_pts + pt) * 3;
    if (which & CAT_OUT_ORIGINAL) {
      T p[3];
      cat_shape_point<T, M>(c.A, c.sp[0], s, last, p);
      o[0] = p[0];
      o[1] = p[1];
      o[2] = p[2];
      o += blk;
    }
    if (need_rot) {
      T p[3];
      cat_shape_point<T, M>(c.A, c.sp[1], s, last, p);
      if (which & CAT_OUT_THETA_ROTATED) {
        o[0] = p[0];
        o[1] = p[1];
        o[2] = p[2];
        o += blk;
      }
      if (which & (CAT_OUT_THETA_ALIGNED | CAT_OUT_FINAL)) {
        // main_fun.py:96-99: A + R(k_theta, -theta)(pt - A)
        const T v[3] = {p[0] - c.A[0], p[1] - c.A[1], p[2] - c.A[2]};
        const T* k = c.k_theta;
        const T cx = k[1] * v[2] - k[2] * v[1], cy = k[2] * v[0] - k[0] * v[2], cz = k[0] * v[1] - k[1] * v[0];
        const T dot = k[0] * v[0] + k[1] * v[1] + k[2] * v[2];
        const T omc = T(1) - c.ct;
        T q[3];
        q[0] = c.A[0] + (v[0] * c.ct + cx * (-c.st) + k[0] * dot * omc);
        q[1] = c.A[1] + (v[1] * c.ct + cy * (-c.st) + k[1] * dot * omc);
        q[2] = c.A[2] + (v[2] * c.ct + cz * (-c.st) + k[2] * dot * omc);
        if (which & CAT_OUT_THETA_ALIGNED) {
          o[0] = q[0];
          o[1] = q[1];
          o[2] = q[2];
          o += blk;
        }
        if (which & CAT_OUT_FINAL) {
          // main_fun.py:106-109: A + R(k_gamma, gamma)(pt - A)
          const T w[3] = {q[0] - c.A[0], q[1] - c.A[1], q[2] - c.A[2]};
          const T* g = c.k_gamma;
          const T gx = g[1] * w[2] - g[2] * w[1], gy = g[2] * w[0] - g[0] * w[2], gz = g[0] * w[1] - g[1] * w[0];
          const T gd = g[0] * w[0] + g[1] * w[1] + g[2] * w[2];
          const T omg = T(1) - c.cg;
          o[0] = c.A[0] + (w[0] * c.cg + gx * c.sg + g[0] * gd * omg);
          o[1] = c.A[1] + (w[1] * c.cg + gy * c.sg + g[1] * gd * omg);
          o[2] = c.A[2] + (w[2] * c.cg + gz * c.sg + g[2] * gd * omg);
        }
      }
    }
  }
}

// np.argmin(curve[:,2]) per curve -- fully_augmented_catenary.py:20-22.  One warp per curve.
template <typename T>
__global__ void __launch_bounds__(256) k_lowest_point(long long n, int n_pts, const T* __restrict__ curves,
                                                      long long* __restrict__ idx_out, T* __restrict__ pt_out) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n) return;
  const T* c = curves + w * (long long)n_pts * 3;
  double best = 0;
  long long bi = 0x7fffffffffffffffLL;
  bool have = false;
  for (int p = lane; p < n_pts; p += 32) {
    const double z = (double)c[3 * p + 2];
    if (!have || cat_better(z, p, best, bi)) {
      best = z;
      bi = p;
      have = true;
    }
  }
  if (!have) best = __longlong_as_double(0x7ff0000000000000LL);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const double ob = __shfl_down_sync(0xffffffffu, best, off);
    const long long oi = __shfl_down_sync(0xffffffffu, bi, off);
    if (cat_better(ob, oi, best, bi)) {
      best = ob;
      bi = oi;
    }
  }
  if (lane == 0) {
    idx_out[w] = bi;
    pt_out[3 * w] = c[3 * bi];
    pt_out[3 * w + 1] = c[3 * bi + 1];
    pt_out[3 * w + 2] = c[3 * bi + 2];
  }
}

// rodrigues_rotation (main_fun.py:18-35) batched: out[i] = R(axis[i*as], angle[i*gs]) v[i].
// axis_stride / angle_stride 0 broadcast one axis / one angle over all vectors (the per-point loops
// of transform_catenary, main_fun.py:96-99,106-109).  origin (3 values, may be NULL) is subtracted
// before and added after the rotation (A + R (pt - A)).
template <typename T>
__global__ void __launch_bounds__(256) k_rodrigues(long long n, const T* __restrict__ v, const T* __restrict__ axis,
                                                   long long axis_stride, const T* __restrict__ angle,
                                                   long long angle_stride, const T* __restrict__ origin,
                                                   T* __restrict__ out) {
  typedef CatMath<T> M;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const T o[3] = {origin ? origin[0] : T(0), origin ? origin[1] : T(0), origin ? origin[2] : T(0)};
  const T vv[3] = {v[3 * i] - o[0], v[3 * i + 1] - o[1], v[3 * i + 2] - o[2]};
  const T* ax = axis + 3 * i * axis_stride;
  const T k[3] = {ax[0], ax[1], ax[2]};
  T s, c;
  M::sincos_(angle[i * angle_stride], &s, &c);
  T r[3];
  cat_rodrigues<T, M>(vv, k, c, s, r);
  out[3 * i] = o[0] + r[0];
  out[3 * i + 1] = o[1] + r[1];
  out[3 * i + 2] = o[2] + r[2];
}

// Span terms used by Catenary.__call__'s scalar outputs and by the tension feature:
// C (main_fun.py:423), horizontal offset of the lowest point from A, lowest z relative to zA
// (SURVEY.md 8a), tension w l / (2 sinh(C l / 2)) with the nan -> w l / 2 fallback (main_fun.py:302-305).
template <typename T>
__global__ void __launch_bounds__(256) k_span_terms(long long n, const T* __restrict__ l, const T* __restrict__ dH,
                                                    const T* __restrict__ L, long long L_stride, T c_lo, T c_hi,
                                                    T w, T* __restrict__ C_out, T* __restrict__ x_low,
                                                    T* __restrict__ z_low, T* __restrict__ tension) {
  typedef CatMath<T> M;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const T li = l[i], hi = dH[i], Li = L[i * L_stride];
  const CatSpan<T> s = cat_span<T, M>(li, hi, Li, c_lo, c_hi, cat_cold<T>());
  if (C_out) C_out[i] = s.C;
  const bool ok = s.C == s.C;
  if (x_low) {
    // l/2 - atanh(dH/L)/C = (l / 2u) (u - atanh(dH/L))
    const T at = T(0.5) * M::log_(M::div_(Li + hi, Li - hi));
    x_low[i] = ok ? M::div_(li * (s.u - at), T(2) * s.u) : M::nan_();
  }
  if (z_low) z_low[i] = cat_z_low<T, M>(s, li, Li, T(0), hi);
  if (tension) tension[i] = ok ? M::div_(w * li, T(2) * s.sh) : w * li * T(0.5);
}
