// cat_kernels_feat.cuh -- the reference's smoothed / second-order feature layouts (second half;
// sm_100a): build_theta_features_valid / build_gamma_features_valid (main_fun.py:504-642), compute_derivatives
// (main_fun.py:645-655), preprocess_signals (main_fun.py:768-776), features_dd (main_fun.py:811-871).
//
// Signals are stored one after the other ([n_sig][len], "column-major" for a feature table), so that the three
// primitives -- FIR with SciPy's 'reflect' boundary (gaussian_filter1d), Savitzky-Golay with SciPy's polynomial edge
// fit (savgol_filter, mode='interp'), np.gradient -- run over all columns of a table in one launch with coalesced
// accesses along time; one more kernel per layout forms the per-row terms and another writes the row-major table.
// Filter taps and the edge-fit matrices are computed on the host (features.py) and passed in: the kernels are generic
// in window length, polynomial order and sigma.  HBM-bound, a few bytes per sample; tables of 1e5..1e7 rows.
#pragma once
#include "cat_math.cuh"

// out[s][i] = sum_j taps[j] * x[s][reflect(i + j - r)],  r = (n_taps - 1) / 2;  reflect = scipy.ndimage 'reflect'
// (half-sample symmetric: d c b a | a b c d | d c b a; any signal length >= 1)
template <typename T>
__global__ void __launch_bounds__(256) k_fir_reflect(long long n_sig, long long len, const T* __restrict__ x,
                                                     int n_taps, const T* __restrict__ taps, T* __restrict__ out) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_sig * len) return;
  const long long s = g / len, i = g - s * len;
  const T* xs = x + s * len;
  const int r = (n_taps - 1) / 2;
  const long long per = 2 * len;
  T acc = T(0);
  for (int j = 0; j < n_taps; ++j) {
    long long k = i + j - r;
    if (k < 0 || k >= len) {
      k %= per;
      if (k < 0) k += per;
      if (k >= len) k = per - 1 - k;
    }
    acc += taps[j] * xs[k];
  }
  out[g] = acc;
}

// scipy.signal.savgol_filter(x, window, order) with mode='interp': interior samples = taps . x[i-h .. i+h]; the first /
// last h = window / 2 samples = edge_l / edge_r (h x window) applied to the first / last `window` samples (the
// least-squares polynomial of that window evaluated in place).  len >= window.
template <typename T>
__global__ void __launch_bounds__(256) k_savgol(long long n_sig, long long len, const T* __restrict__ x, int window,
                                                const T* __restrict__ taps, const T* __restrict__ edge_l,
                                                const T* __restrict__ edge_r, T* __restrict__ out) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_sig * len) return;
  const long long s = g / len, i = g - s * len;
  const T* xs = x + s * len;
  const int h = window / 2;
  const T* w;
  const T* src;
  if (i < h) {
    w = edge_l + i * window;
    src = xs;
  } else if (i >= len - h) {
    w = edge_r + (i - (len - h)) * window;
    src = xs + (len - window);
  } else {
    w = taps;
    src = xs + (i - h);
  }
  T acc = T(0);
  for (int j = 0; j < window; ++j) acc += w[j] * src[j];
  out[g] = acc;
}

// np.gradient(y, t), edge_order = 1: NumPy takes (f[i+1] - f[i-1]) / (2 h) when every spacing is bitwise equal
// (uniform != 0, decided on the host exactly as NumPy does) and the second-order non-uniform formula otherwise;
// one-sided first differences at both ends.  len >= 2.
template <typename T>
__device__ __forceinline__ T cat_gradient_at(const T* __restrict__ y, const T* __restrict__ t, long long i,
                                             long long len, int uniform) {
  if (i == 0) return (y[1] - y[0]) / (t[1] - t[0]);
  if (i == len - 1) return (y[i] - y[i - 1]) / (t[i] - t[i - 1]);
  if (uniform) return (y[i + 1] - y[i - 1]) / (T(2) * (t[1] - t[0]));
  const T hs = t[i] - t[i - 1], hd = t[i + 1] - t[i];
  const T a = -hd / (hs * (hd + hs)), b = (hd - hs) / (hd * hs), c = hs / (hd * (hd + hs));
  return a * y[i - 1] + b * y[i] + c * y[i + 1];
}

template <typename T>
__global__ void __launch_bounds__(256) k_gradient(long long n_sig, long long len, const T* __restrict__ y,
                                                  const T* __restrict__ t, int uniform, T* __restrict__ out) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_sig * len) return;
  const long long s = g / len, i = g - s * len;
  out[g] = cat_gradient_at<T>(y + s * len, t, i, len, uniform);
}

// per-row cable-relative terms shared by the layouts below: l = |rel_xy|, v_surge = V . unit_rel,
// v_sway = |V x unit_rel| with unit_rel = rel / (|rel| + 1e-8) (main_fun.py:523-529, 843-847); V scaled by v_scale
// first (features_dd works in m/s: main_fun.py:815).  Writes signals [l | v_surge | v_sway | Vx | Vy | Vz] of length T.
template <typename T>
__global__ void __launch_bounds__(256) k_cable_terms(long long T_len, const T* __restrict__ P0, const T* __restrict__ P1,
                                                     const T* __restrict__ V1, T v_scale, T* __restrict__ sig) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= T_len) return;
  const T rx = P1[3 * i] - P0[3 * i], ry = P1[3 * i + 1] - P0[3 * i + 1], rz = P1[3 * i + 2] - P0[3 * i + 2];
  const T v[3] = {V1[3 * i] * v_scale, V1[3 * i + 1] * v_scale, V1[3 * i + 2] * v_scale};
  const T l2 = rx * rx + ry * ry;
  const T den = sqrt(l2 + rz * rz) + T(1e-8);
  const T u[3] = {rx / den, ry / den, rz / den};
  const T cx = v[1] * u[2] - v[2] * u[1], cy = v[2] * u[0] - v[0] * u[2], cz = v[0] * u[1] - v[1] * u[0];
  sig[i] = sqrt(l2);
  sig[T_len + i] = v[0] * u[0] + v[1] * u[1] + v[2] * u[2];
  sig[2 * T_len + i] = sqrt(cx * cx + cy * cy + cz * cz);
  sig[3 * T_len + i] = v[0];
  sig[4 * T_len + i] = v[1];
  sig[5 * T_len + i] = v[2];
}

// build_theta_features_valid / build_gamma_features_valid rows (main_fun.py:549-569, 620-640), 13 columns:
//   a x / (l + 1e-8), x / (l + 1e-8), a x, x, x^2 / (l + 1e-8), theta, gamma, b y / (l + 1e-8), y / (l + 1e-8), b y, y,
//   dtheta, dgamma      with (a, x, b, y) = (theta, v_surge, gamma, v_sway) for theta, (gamma, v_sway, theta, v_surge)
// for gamma.  th1 / ga1: the (T,1)-filtered angles (= raw x weight sum, see features.py); dth / dga: np.gradient of the
// properly smoothed ones.
template <typename T>
__global__ void __launch_bounds__(256) k_valid_rows(long long T_len, const T* __restrict__ sig,
                                                    const T* __restrict__ th1, const T* __restrict__ ga1,
                                                    const T* __restrict__ dth, const T* __restrict__ dga,
                                                    int for_gamma, T* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= T_len) return;
  const T l = sig[i] + T(1e-8), vs = sig[T_len + i], vw = sig[2 * T_len + i];
  const T th = th1[i], ga = ga1[i];
  const T a = for_gamma ? ga : th, x = for_gamma ? vw : vs, b = for_gamma ? th : ga, y = for_gamma ? vs : vw;
  T* o = out + i * 13;
  o[0] = a * x / l;
  o[1] = x / l;
  o[2] = a * x;
  o[3] = x;
  o[4] = x * x / l;
  o[5] = th;
  o[6] = ga;
  o[7] = b * y / l;
  o[8] = y / l;
  o[9] = b * y;
  o[10] = y;
  o[11] = dth[i];
  o[12] = dga[i];
}

// features_dd rows (main_fun.py:852-873): sm = [theta_s | gamma_s] (Savitzky-Golay), dsm = np.gradient(sm),
// d2 = np.gradient(dsm), dsig = np.gradient of sig's [v_surge | v_sway | Vx | Vy | Vz]; sig as k_cable_terms wrote it.
//   features (T,14): theta_s, gamma_s, theta_dot, gamma_dot, v_sway, v_surge, a_sway, a_surge, V (3), a (3)
//   targets (T,2): theta_ddot, gamma_ddot
template <typename T>
__global__ void __launch_bounds__(256) k_dd_rows(long long T_len, const T* __restrict__ sig, const T* __restrict__ sm,
                                                 const T* __restrict__ dsm, const T* __restrict__ dsig,
                                                 const T* __restrict__ d2, T* __restrict__ feats,
                                                 T* __restrict__ targets) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= T_len) return;
  T* o = feats + i * 14;
  o[0] = sm[i];
  o[1] = sm[T_len + i];
  o[2] = dsm[i];
  o[3] = dsm[T_len + i];
  o[4] = sig[2 * T_len + i];    // v_sway
  o[5] = sig[T_len + i];        // v_surge
  o[6] = dsig[T_len + i];       // a_sway
  o[7] = dsig[i];               // a_surge
  o[8] = sig[3 * T_len + i];
  o[9] = sig[4 * T_len + i];
  o[10] = sig[5 * T_len + i];
  o[11] = dsig[2 * T_len + i];
  o[12] = dsig[3 * T_len + i];
  o[13] = dsig[4 * T_len + i];
  targets[2 * i] = d2[i];
  targets[2 * i + 1] = d2[T_len + i];
}
