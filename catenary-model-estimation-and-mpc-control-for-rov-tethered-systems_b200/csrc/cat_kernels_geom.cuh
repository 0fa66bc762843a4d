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
  const CatSpan<T> s = cat_span_ldh<T, M>(l[i], dH[i], L[i * L_stride], c_lo, c_hi);   // fp32: N formed in fp64
  C_out[i] = s.C;
}

// ------------------------------------------------------------------------------------------
// Curve shapes: Catenary.__call__ (pympc; builder-defined sampling, see oracle.catenary_points)
// fused with transform_catenary (main_fun.py:38-111).
//
// Block = CAT_SHAPE_TILE configurations.
// Phase 1: thread t solves configuration t of the tile (span A->B and, for the rotated outputs, the
//   span A->B' of main_fun.py:92-93) and parks in shared memory, per span, the constants of the sag
//   profile and, per requested output, an affine map: every curve lies in the vertical plane through
//   its chord, so a sample is  A + s a + zoff(s) b  with  a = R (rel_x, rel_y, 0),  b = R e_z  and R the
//   composition of the Rodrigues rotations of main_fun.py:96-99,106-109 (identity for the first two
//   outputs).  The two per-point rotations of the reference collapse into 6 FMAs per point.
// Phase 2: one warp per configuration, lanes over points.  zoff(s) = (cosh(2u s - v_low) - cosh v_low)/C
//   is evaluated with the polynomial sinh/cosh once per lane and advanced by 32 samples with the
//   addition theorem (4 FMAs), so a 100-point curve costs 32 transcendental evaluations, not 100.
//   A warp writes 32 consecutive (x,y,z) triples = 768 contiguous bytes of the AoS (n, n_pts, 3) output.
// ------------------------------------------------------------------------------------------
#define CAT_SHAPE_TILE 128
#ifndef CAT_SHAPE_MINB
#define CAT_SHAPE_MINB 5    // resident blocks per SM the register budget is cut for
#endif
// lanes per configuration in phase 2 (power of two).  Measured on B200, 1,048,576 x 100 points, final curve:
//   fp64: 32 lanes 0.53 ms, 16 lanes 0.57 ms, 8 lanes 0.91 ms -- the stores want one contiguous 768-byte run per
//         warp instruction; splitting the warp over 2 or 4 curves costs more in the memory system than the saved
//         cosh/sinh evaluations and idle lanes give back
//   fp32: 16 lanes 0.28 ms vs 32 lanes 0.33 ms (each half-warp still writes a 192-byte run)
template <typename T> struct CatShapeGroup { static constexpr int lanes = sizeof(T) == 4 ? 16 : 32; };
// Stores stay scalar (three 8- or 4-byte stores per lane at a 24- or 12-byte stride).  Round 2 measured the
// alternative -- park the warp's samples in shared memory and write each 32-point hop as consecutive float4 / double2
// -- on one B200 (profiles/r2_ab_shape_vec_stores.txt): fp64 final curve 0.744 vs 0.449 ms, fp32 0.469 vs 0.268 ms.
// The two warp barriers and the staging traffic per hop cost far more than the narrower stores save; L2 already
// merges the strided stores into full sectors (DRAM write bytes == algorithmic bytes in the ncu capture).
// A second form without staging -- each lane owns two consecutive points (the second by a one-sample shift of
// cosh/sinh) and writes its 48 / 24 bytes as three vectors, half the hops -- lost as well
// (profiles/r2_ab_shape_pair_stores.txt: fp64 0.618 vs 0.447 ms, fp32 0.296 vs 0.268 ms): every vector store then
// covers 16 of each 48 bytes of a 1.5 KB window, all sectors partial, and the L2 write path is the bound.
// What does pay, for the masks that write two or more rotated curves (their stores interleave inside one span loop), is
// TWO stores per point -- a 2-value vector + a scalar, chosen by the point's parity (template parameter ST2; the host
// picks it, profiles/r2_ab_shape_two_stores.txt): all four curves 1.87 -> 1.57 ms in fp64, 1.34 -> 0.98 ms in fp32.
#define CAT_SHAPE_RESEED 16    // (power of two) re-evaluate cosh/sinh directly every 16 hops (bounds the recurrence drift)

template <typename T> struct CatShapeSpan {
  T two_u;      // C * l
  T v_low;      // C * x_low = u - atanh(dH/L)
  T inv_C;      // 1/C
  T cm_low;     // cosh(v_low) - 1
  T cmD, shD;   // cosh(D) - 1, sinh(D) for the 32-sample hop D = 2u * 32/(n_pts-1)
  int valid;
};

template <typename T> struct CatShapeMap {
  T a[3], b[3];   // sample = A + s a + zoff b
};

// Shared-memory record of one configuration, sized by what `which` asks for (the full record is 352 B in fp64,
// the final-curve-only record 144 B), as PAIRS of T so that phase 2 reads it with 16-byte LDS (fp64):
//   (A0,A1) (A2,-) | per needed span: (two_u,v_low) (inv_C,cm_low) (cmD,shD) (valid,-) | per requested map:
//   (a0,a1) (a2,b0) (b1,b2).  Phase-2 reads are warp-uniform (broadcast), so only alignment matters there.
#define CAT_SHAPE_SPAN_W 4   // pairs
#define CAT_SHAPE_MAP_W 3    // pairs
template <typename T> struct alignas(2 * sizeof(T)) CatPair {
  T x, y;
};
__host__ __device__ __forceinline__ int cat_shape_record_len(int which) {   // in pairs
  const int nspan = ((which & CAT_OUT_ORIGINAL) ? 1 : 0) + ((which & ~CAT_OUT_ORIGINAL & 15) ? 1 : 0);
  int nmap = 0;
  for (int m = 0; m < 4; ++m) nmap += (which >> m) & 1;
  return 2 + CAT_SHAPE_SPAN_W * nspan + CAT_SHAPE_MAP_W * nmap;
}

template <typename T>
__device__ __forceinline__ void cat_shape_store_span(CatPair<T>* r, const CatShapeSpan<T>& sp) {
  r[0] = CatPair<T>{sp.two_u, sp.v_low};
  r[1] = CatPair<T>{sp.inv_C, sp.cm_low};
  r[2] = CatPair<T>{sp.cmD, sp.shD};
  r[3] = CatPair<T>{T(sp.valid), T(0)};
}

// rel_exact (fp32 only, may be null): B - A formed in fp64 from the fp32 end points, so that N = L^2 - dH^2 - l^2
// carries one rounding instead of the eps L^2 / N of the fp32 subtractions (cat_span_prep_exact)
template <typename T, typename M>
__device__ __forceinline__ void cat_shape_span_setup(const T rel[3], const double* rel_exact, T L, T c_lo, T c_hi,
                                                     T hop, CatShapeSpan<T>& sp, T* C_out) {
  const T l = M::sqrt_(rel[0] * rel[0] + rel[1] * rel[1]);
  const T dH = rel[2];
  CatSpan<T> s;
  if constexpr (sizeof(T) == 4) {
    if (rel_exact) {
      const double l2 = rel_exact[0] * rel_exact[0] + rel_exact[1] * rel_exact[1];
      s = cat_span_from<T, M>(cat_span_prep_exact<M>(l, l2, rel_exact[2], (double)L), l, dH, L, c_lo, c_hi,
                              cat_cold<T>());
    } else {
      s = cat_span<T, M>(l, dH, L, c_lo, c_hi, cat_cold<T>());
    }
  } else {
    s = cat_span<T, M>(l, dH, L, c_lo, c_hi, cat_cold<T>());
  }
  sp.valid = (s.C == s.C) ? 1 : 0;
  if (C_out) *C_out = s.C;
  if (sp.valid) {
    sp.two_u = T(2) * s.u;
    sp.v_low = s.u - T(0.5) * M::log_(M::div_(L + dH, L - dH));
    sp.inv_C = M::rcp_(s.C);
    T sm, sh;
    cat_hyp<T, M>(M::abs_(sp.v_low), sm, sp.cm_low, sh);
    cat_hyp<T, M>(sp.two_u * hop, sm, sp.cmD, sp.shD);
  } else {
    sp.two_u = T(0);
    sp.v_low = T(0);
    sp.inv_C = T(0);
    sp.cm_low = T(0);
    sp.cmD = T(0);
    sp.shD = T(0);
  }
}

// R = c I + s [k]x + (1-c) k k^T applied to v (k unit), main_fun.py:31-35
template <typename T>
__device__ __forceinline__ void cat_rot_apply(const T k[3], T c, T s, const T v[3], T out[3]) {
  const T cx = k[1] * v[2] - k[2] * v[1];
  const T cy = k[2] * v[0] - k[0] * v[2];
  const T cz = k[0] * v[1] - k[1] * v[0];
  const T dot = k[0] * v[0] + k[1] * v[1] + k[2] * v[2];
  const T omc = T(1) - c;
  out[0] = v[0] * c + cx * s + k[0] * dot * omc;
  out[1] = v[1] * c + cy * s + k[1] * dot * omc;
  out[2] = v[2] * c + cz * s + k[2] * dot * omc;
}

template <typename T>
__device__ __forceinline__ void cat_shape_map_set(CatShapeMap<T>& m, const T a[3], const T b[3]) {
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    m.a[j] = a[j];
    m.b[j] = b[j];
  }
}

// W = the output mask as a compile-time constant (0: taken from which_rt at run time).  With W fixed the record
// layout, the per-map tests of the point loop and the map loads are resolved at compile time: the maps of a
// configuration sit in registers across its points (profiles/r1_ncu_full_k_shape_f64_v3c.csv: the run-time form
// spends 55 instructions per 32-point hop, of which ~20 are mask tests, offset arithmetic and repeated LDS).
// ST2: two stores per point (one 2-value vector + one scalar) instead of three scalars -- see the store site.
template <typename T, int W, bool ST2 = false>
__global__ void __launch_bounds__(CAT_SHAPE_TILE, CAT_SHAPE_MINB) k_shape(long long n, const T* __restrict__ A,
                                                          const T* __restrict__ B, const T* __restrict__ L,
                                                          long long L_stride, const T* __restrict__ theta,
                                                          const T* __restrict__ gamma, int n_pts, int which_rt,
                                                          T c_lo, T c_hi, T* __restrict__ out,
                                                          T* __restrict__ C_out, unsigned char* __restrict__ valid) {
  typedef CatMath<T> M;
  const int which = W ? W : which_rt;
  constexpr bool HOIST = W != 0 && ((W & 1) + ((W >> 1) & 1) + ((W >> 2) & 1) + ((W >> 3) & 1)) <= 2;
  extern __shared__ __align__(16) unsigned char cat_shape_smem[];
  CatPair<T>* const rec = reinterpret_cast<CatPair<T>*>(cat_shape_smem);
  const int rec_len = cat_shape_record_len(which);
  const bool need0 = (which & CAT_OUT_ORIGINAL) != 0;
  const int span1_at = 2 + (need0 ? CAT_SHAPE_SPAN_W : 0);
  const int maps_at = span1_at + ((which & ~CAT_OUT_ORIGINAL & 15) ? CAT_SHAPE_SPAN_W : 0);
  const long long base = (long long)blockIdx.x * CAT_SHAPE_TILE;
  const int tid = threadIdx.x;
  const long long i = base + tid;
  const bool need_rot = (which & (CAT_OUT_THETA_ROTATED | CAT_OUT_THETA_ALIGNED | CAT_OUT_FINAL)) != 0;
  const T inv_nm1 = (n_pts > 1) ? M::div_(T(1), T(n_pts - 1)) : T(0);
  const T hop = T(CatShapeGroup<T>::lanes) * inv_nm1;
  if (i < n) {
    struct {
      CatShapeSpan<T> sp[2];
      CatShapeMap<T> map[4];
    } c;
    CatPair<T>* const r = rec + tid * rec_len;
    const T a[3] = {A[3 * i], A[3 * i + 1], A[3 * i + 2]};
    const T b[3] = {B[3 * i], B[3 * i + 1], B[3 * i + 2]};
    const T Li = L[i * L_stride];
    r[0] = CatPair<T>{a[0], a[1]};
    r[1] = CatPair<T>{a[2], T(0)};
    const T conn[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]};
    const T ez[3] = {T(0), T(0), T(1)};
    const T zero3[3] = {T(0), T(0), T(0)};
    T Cv;
    const double conn_d[3] = {(double)b[0] - (double)a[0], (double)b[1] - (double)a[1], (double)b[2] - (double)a[2]};
    cat_shape_span_setup<T, M>(conn, sizeof(T) == 4 ? conn_d : nullptr, Li, c_lo, c_hi, hop, c.sp[0], &Cv);
    if (C_out) C_out[i] = Cv;
    int vbits = c.sp[0].valid;
    if (c.sp[0].valid) {
      const T axy[3] = {conn[0], conn[1], T(0)};
      cat_shape_map_set<T>(c.map[0], axy, ez);
    } else {                                   // straight chord [A, B] (main_fun.py:67-69)
      cat_shape_map_set<T>(c.map[0], conn, zero3);
    }
    if (need_rot) {
      const T th = theta ? theta[i] : T(0);
      const T ga = gamma ? gamma[i] : T(0);
      T st, ct, sg, cg;
      M::sincos_(th, &st, &ct);
      M::sincos_(ga, &sg, &cg);
      T kt[3];
      cat_theta_axis<T, M>(conn, kt);
      // main_fun.py:92  rotated_B = A + rodrigues(conn, k_theta, theta)
      T rb[3];
      cat_rodrigues<T, M>(conn, kt, ct, st, rb);
      const T bq[3] = {a[0] + rb[0], a[1] + rb[1], a[2] + rb[2]};
      const T rel1[3] = {bq[0] - a[0], bq[1] - a[1], bq[2] - a[2]};
      cat_shape_span_setup<T, M>(rel1, nullptr, Li, c_lo, c_hi, hop, c.sp[1], (T*)nullptr);
      vbits |= c.sp[1].valid << 1;
      // axes as rodrigues_rotation normalises them (main_fun.py:30)
      const T nt = M::sqrt_(kt[0] * kt[0] + kt[1] * kt[1] + kt[2] * kt[2]);
      const T kth[3] = {M::div_(kt[0], nt), M::div_(kt[1], nt), M::div_(kt[2], nt)};
      // main_fun.py:102-103 gamma axis = conn/|conn| (then normalised again inside rodrigues)
      const T nc = M::sqrt_(conn[0] * conn[0] + conn[1] * conn[1] + conn[2] * conn[2]);
      const T kg0[3] = {M::div_(conn[0], nc), M::div_(conn[1], nc), M::div_(conn[2], nc)};
      const T ng = M::sqrt_(kg0[0] * kg0[0] + kg0[1] * kg0[1] + kg0[2] * kg0[2]);
      const T kga[3] = {M::div_(kg0[0], ng), M::div_(kg0[1], ng), M::div_(kg0[2], ng)};
      // in-plane basis of the rotated span: (a1, b1); chord fallback: (rel1, 0)
      T a1[3], b1[3];
      if (c.sp[1].valid) {
        a1[0] = rel1[0];
        a1[1] = rel1[1];
        a1[2] = T(0);
        b1[0] = T(0);
        b1[1] = T(0);
        b1[2] = T(1);
      } else {
        a1[0] = rel1[0];
        a1[1] = rel1[1];
        a1[2] = rel1[2];
        b1[0] = b1[1] = b1[2] = T(0);
      }
      cat_shape_map_set<T>(c.map[1], a1, b1);
      // main_fun.py:96-99: A + R(k_theta, -theta)(pt - A)
      T a2[3], b2[3];
      cat_rot_apply<T>(kth, ct, -st, a1, a2);
      cat_rot_apply<T>(kth, ct, -st, b1, b2);
      cat_shape_map_set<T>(c.map[2], a2, b2);
      // main_fun.py:106-109: A + R(k_gamma, gamma)(pt - A)
      T a3[3], b3[3];
      cat_rot_apply<T>(kga, cg, sg, a2, a3);
      cat_rot_apply<T>(kga, cg, sg, b2, b3);
      cat_shape_map_set<T>(c.map[3], a3, b3);
    }
    if (valid) valid[i] = (unsigned char)vbits;
    if (need0) cat_shape_store_span<T>(r + 2, c.sp[0]);
    if (need_rot) cat_shape_store_span<T>(r + span1_at, c.sp[1]);
    int slot = 0;
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      if (!(which & (1 << m))) continue;
      CatPair<T>* rm = r + maps_at + CAT_SHAPE_MAP_W * slot;
      rm[0] = CatPair<T>{c.map[m].a[0], c.map[m].a[1]};
      rm[1] = CatPair<T>{c.map[m].a[2], c.map[m].b[0]};
      rm[2] = CatPair<T>{c.map[m].b[1], c.map[m].b[2]};
      ++slot;
    }
  }
  __syncthreads();

  const int n_here = (int)((n - base < CAT_SHAPE_TILE) ? (n - base) : CAT_SHAPE_TILE);
  const int lane = tid & 31, warp = tid >> 5, n_warps = CAT_SHAPE_TILE / 32;
  const long long blk = n * (long long)n_pts * 3;   // elements per output block
  // output-block offset and record offset of each requested map (slot = rank of m among the requested ones)
  long long m_out[4];
  int m_rec[4];
  {
    int slot = 0;
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      m_out[m] = slot * blk;
      m_rec[m] = maps_at + CAT_SHAPE_MAP_W * slot;
      slot += (which >> m) & 1;
    }
  }
  // G lanes walk one configuration, so a warp serves 32/G configurations at once (G = 32 in fp64, 16 in fp32)
  constexpr int G = CatShapeGroup<T>::lanes;
  const int grp = lane / G, gl = lane % G;
  const T lane_f = T(gl);
  const int n_hops = (n_pts + G - 1) / G;
  // ST2 (host: even n_pts, `out` aligned to two values): G is even, so a lane's points all have the lane's parity,
  // every curve and every output block starts on a 2-value boundary
  const bool odd = (gl & 1) != 0;
  for (int c0 = warp * (32 / G); c0 < n_here; c0 += n_warps * (32 / G)) {
    const bool have = c0 + grp < n_here;
    const int ci = have ? c0 + grp : n_here - 1;   // idle groups shadow the last configuration and do not store
    const CatPair<T>* const r = rec + ci * rec_len;
    const CatPair<T> rA = r[0];
    const T A0 = rA.x, A1 = rA.y, A2 = r[1].x;
    T* const o_lane = out + ((base + ci) * (long long)n_pts + gl) * 3;
#pragma unroll
    for (int spi = 0; spi < 2; ++spi) {
      const int outs = spi == 0 ? (which & CAT_OUT_ORIGINAL) : (which & ~CAT_OUT_ORIGINAL & 15);
      if (!outs) continue;
      const CatPair<T>* const rs = r + (spi == 0 ? 2 : span1_at);
      const CatPair<T> s0 = rs[0], s1 = rs[1], s2 = rs[2];
      const T two_u = s0.x, v_low = s0.y, inv_C = s1.x, cm_low = s1.y, cmD = s2.x, shD = s2.y;
      const bool ok = rs[3].x != T(0);
      CatPair<T> mq[4][3];     // maps of this span's outputs, held in registers when the mask is a constant
      if constexpr (HOIST) {
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          if ((m == 0) != (spi == 0) || !(outs & (1 << m))) continue;
          const CatPair<T>* mp = r + m_rec[m];
          mq[m][0] = mp[0];
          mq[m][1] = mp[1];
          mq[m][2] = mp[2];
        }
      }
      T cm = T(0), sh = T(0);
      T ptf = lane_f;          // T(pt), stepped exactly
      T* o = o_lane;           // &out[cfg][pt][0] of the first requested block, stepped by G points
      // uniform trip count (idle lanes of the last hop compute and do not store): no divergent loop exit
      for (int hop = 0; hop < n_hops; ++hop, ptf += T(G), o += 3 * G) {
        const int pt = gl + G * hop;
        const T s = (pt == n_pts - 1) ? T(1) : ptf * inv_nm1;
        if ((hop & (CAT_SHAPE_RESEED - 1)) == 0) {   // warp-uniform
          const T t = two_u * s - v_low;
          T sm;
          cat_hyp<T, M>(M::abs_(t), sm, cm, sh);
          sh = t < T(0) ? -sh : sh;
        } else {
          // cosh(t + D) - 1 = cm + cm cmD + cmD + sh shD ;  sinh(t + D) = sh + sh cmD + (cm + 1) shD
          const T cn = M::fma_(cm, cmD, cm) + M::fma_(sh, shD, cmD);
          const T sn = M::fma_(sh, cmD, sh) + M::fma_(cm, shD, shD);
          cm = cn;
          sh = sn;
        }
        const T zoff = ok ? (cm - cm_low) * inv_C : T(0);   // chord fallback: spans outside the solve domain
        if (pt >= n_pts || !have) continue;
        // sample = A + s a + zoff b as two FMAs per coordinate
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          if ((m == 0) != (spi == 0) || !(outs & (1 << m))) continue;
          const CatPair<T>* mp = r + m_rec[m];   // (a0,a1) (a2,b0) (b1,b2)
          CatPair<T> p0, p1, p2;
          if constexpr (HOIST) {
            p0 = mq[m][0];
            p1 = mq[m][1];
            p2 = mq[m][2];
          } else {
            p0 = mp[0];
            p1 = mp[1];
            p2 = mp[2];
          }
          T* om = o + m_out[m];
          const T X = M::fma_(zoff, p1.y, M::fma_(s, p0.x, A0));
          const T Y = M::fma_(zoff, p2.x, M::fma_(s, p0.y, A1));
          const T Z = M::fma_(zoff, p2.y, M::fma_(s, p1.x, A2));
          if constexpr (ST2) {
            // two stores per point instead of three: the point's 3 values start on a 2-value boundary when its index
            // is even ((x,y) as one vector, then z) and one value past it when odd (x, then (y,z) as one vector)
            *reinterpret_cast<CatPair<T>*>(om + (odd ? 1 : 0)) = CatPair<T>{odd ? Y : X, odd ? Z : Y};
            om[odd ? 0 : 2] = odd ? X : Z;
          } else {
            om[0] = X;
            om[1] = Y;
            om[2] = Z;
          }
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

// ------------------------------------------------------------------------------------------
// compute_catenary_3D (models/catenary_3d.py:5-39), batched: the reference's alternative sampler --
// taut rope -> np.linspace(p0, p1); else the fixed point a <- a L / (2 a sinh(d / 2a)) from a = d/2
// (at most 100 passes, stop when |a_new - a| < 1e-6) and the symmetric sag profile
// sag(t) = a cosh(d/2a) - a cosh((t d - d/2)/a) subtracted from the chord's z.
// k_catenary3d_solve: one thread per rope -> a (-1 marks the taut case);  k_catenary3d_points: one thread
// per sample, AoS (n, n_pts, 3) output so a warp writes 768 contiguous bytes.
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_catenary3d_solve(long long n, const T* __restrict__ p0,
                                                          const T* __restrict__ p1, const T* __restrict__ L,
                                                          long long L_stride, T* __restrict__ a_out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const T dx = p1[3 * i] - p0[3 * i], dy = p1[3 * i + 1] - p0[3 * i + 1], dz = p1[3 * i + 2] - p0[3 * i + 2];
  const T dist = sqrt(dx * dx + dy * dy + dz * dz);
  const T rope = L[i * L_stride];
  if (rope <= dist) {
    a_out[i] = T(-1);   // taut marker (a itself can legitimately end at 0, inf or NaN: the reference's map is repelling)
    return;
  }
  T a = dist / T(2);
  for (int it = 0; it < 100; ++it) {
    const T lhs = T(2) * a * sinh(dist / (T(2) * a));
    const T a_new = a * rope / lhs;
    const bool done = fabs(a_new - a) < T(1e-6);
    a = a_new;
    if (done) break;
  }
  a_out[i] = a;
}

template <typename T>
__global__ void __launch_bounds__(256) k_catenary3d_points(long long n, int n_pts, const T* __restrict__ p0,
                                                           const T* __restrict__ p1, const T* __restrict__ a_in,
                                                           T* __restrict__ out) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n * (long long)n_pts) return;
  const long long c = g / n_pts;
  const int i = (int)(g - c * n_pts);
  const T x0 = p0[3 * c], y0 = p0[3 * c + 1], z0 = p0[3 * c + 2];
  const T dx = p1[3 * c] - x0, dy = p1[3 * c + 1] - y0, dz = p1[3 * c + 2] - z0;
  const T a = a_in[c];
  T* o = out + g * 3;
  if (a == T(-1)) {                      // taut: np.linspace(p0, p1, n): start + i * (delta / (n-1)), last = stop
    if (i == n_pts - 1 && n_pts > 1) {
      o[0] = p1[3 * c];
      o[1] = p1[3 * c + 1];
      o[2] = p1[3 * c + 2];
    } else {
      const T div = T(n_pts > 1 ? n_pts - 1 : 1);
      o[0] = x0 + T(i) * (dx / div);
      o[1] = y0 + T(i) * (dy / div);
      o[2] = z0 + T(i) * (dz / div);
    }
    return;
  }
  const T dist = sqrt(dx * dx + dy * dy + dz * dz);
  const T half = dist / T(2);
  const T off = a * cosh(half / a);
  const T t = T(i) / T(n_pts - 1);
  const T sag = off - a * cosh((t * dist - half) / a);
  o[0] = x0 + dx * t;
  o[1] = y0 + dy * t;
  o[2] = (z0 + dz * t) - sag;
}

// catenary_equation (models/catenary-3d-visualization/src/catenary_model.py:10-12): y = (cosh(C x) - 1) / C,
// C one value (stride 0) or one per sample.  cosh - 1 is formed without cancellation (cat_hyp).
template <typename T>
__global__ void __launch_bounds__(256) k_catenary_equation(long long n, const T* __restrict__ x,
                                                           const T* __restrict__ C, long long C_stride,
                                                           T* __restrict__ y) {
  typedef CatMath<T> M;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const T c = C[i * C_stride];
  T sm, cm, sh;
  cat_hyp<T, M>(M::abs_(c * x[i]), sm, cm, sh);
  y[i] = cm / c;
}
