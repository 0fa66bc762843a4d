// cat_kernels_dyn.cuh -- equation evaluation, feature build and open-loop integrators (sm_100a).
#pragma once
#include "cat_math.cuh"
#include "gen/dyn_eqs.cuh"

template <typename T>
__device__ __forceinline__ void cat_load_row(const T* __restrict__ X, long long row, bool scale, T (&x)[18]) {
#pragma unroll
  for (int j = 0; j < 18; ++j) {
    const T v = X[row * 18 + j];
    x[j] = scale ? (v - CatScaler<T>::mean(j)) * CatScaler<T>::inv_scale(j) : v;
  }
}

// PySRRegressor.predict stand-in: out[i] = eq(X[i,:])   (test_temp.py:53-54)
template <typename T>
__global__ void __launch_bounds__(256) k_predict(long long n, const T* __restrict__ X, int eq_id, int apply_scaler,
                                                 T* __restrict__ out) {
  typedef CatMath<T> M;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  T x[18];
  cat_load_row<T>(X, i, apply_scaler != 0, x);
  out[i] = cat_eval_equation<T, M>(eq_id, x);
}

// Increment of one open-loop step, simulate_rk4_theta_gamma.py:55-66 (method 0) or
// main_fun.py:758-762 (method 1, Euler).  inc[traj][i] for i = 1..T-1; the stage evaluations of
// all steps are independent because the reference never feeds y back into f (k2 == k3).
template <typename T>
__global__ void __launch_bounds__(256) k_rk4_increments(long long n_traj, long long T_len,
                                                        const T* __restrict__ X, const T* __restrict__ time,
                                                        int eq_id, int method, T* __restrict__ inc) {
  typedef CatMath<T> M;
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long per = T_len - 1;
  if (g >= n_traj * per) return;
  const long long tr = g / per;
  const long long i = g - tr * per + 1;
  const T dt = time[i] - time[i - 1];
  T x0[18];
  cat_load_row<T>(X, tr * T_len + i - 1, false, x0);
  const T k1 = cat_eval_equation<T, M>(eq_id, x0);
  T r;
  if (method == 1) {
    r = k1 * dt;
  } else {
    T x1[18], xm[18];
    cat_load_row<T>(X, tr * T_len + i, false, x1);
#pragma unroll
    for (int j = 0; j < 18; ++j) xm[j] = (x0[j] + x1[j]) / T(2);
    const T k2 = cat_eval_equation<T, M>(eq_id, xm);
    const T k3 = k2;
    const T k4 = cat_eval_equation<T, M>(eq_id, x1);
    r = (dt / T(6)) * (k1 + T(2) * k2 + T(2) * k3 + k4);
  }
  inc[tr * T_len + i] = r;
}

// y[i] = y[i-1] + inc[i], strictly in the reference's order (one thread per trajectory).
template <typename T>
__global__ void __launch_bounds__(128) k_rk4_accumulate(long long n_traj, long long T_len, const T* __restrict__ y0,
                                                        const T* __restrict__ inc, T* __restrict__ y) {
  const long long tr = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tr >= n_traj) return;
  const T* in = inc + tr * T_len;
  T* o = y + tr * T_len;
  T acc = y0[tr];
  o[0] = acc;
  long long i = 1;
  for (; i + 8 <= T_len; i += 8) {
    T v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = in[i + j];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc = acc + v[j];
      o[i + j] = acc;
    }
  }
  for (; i < T_len; ++i) {
    acc = acc + in[i];
    o[i] = acc;
  }
}

// extract_features, simple.py:15-41, on columns (T,3)/(T); A1 = np.gradient(V1, time) with numpy's
// second-order non-uniform interior formula and first-order edges.
template <typename T>
__global__ void __launch_bounds__(256) k_features(long long T_len, const T* __restrict__ P0, const T* __restrict__ P1,
                                                  const T* __restrict__ V1, const T* __restrict__ time,
                                                  const T* __restrict__ theta, const T* __restrict__ gamma,
                                                  int apply_scaler, T* __restrict__ X) {
  typedef CatMath<T> M;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= T_len) return;
  T a1[3];
  if (T_len < 2) {
    a1[0] = a1[1] = a1[2] = T(0);
  } else if (i == 0) {
    const T h = time[1] - time[0];
#pragma unroll
    for (int j = 0; j < 3; ++j) a1[j] = (V1[3 + j] - V1[j]) / h;
  } else if (i == T_len - 1) {
    const T h = time[i] - time[i - 1];
#pragma unroll
    for (int j = 0; j < 3; ++j) a1[j] = (V1[3 * i + j] - V1[3 * (i - 1) + j]) / h;
  } else {
    const T hs = time[i] - time[i - 1], hd = time[i + 1] - time[i];
    const T a = -hd / (hs * (hd + hs)), b = (hd - hs) / (hd * hs), c = hs / (hd * (hd + hs));
#pragma unroll
    for (int j = 0; j < 3; ++j) a1[j] = a * V1[3 * (i - 1) + j] + b * V1[3 * i + j] + c * V1[3 * (i + 1) + j];
  }
  const T p0[3] = {P0[3 * i], P0[3 * i + 1], P0[3 * i + 2]};
  const T p1[3] = {P1[3 * i], P1[3 * i + 1], P1[3 * i + 2]};
  const T v1[3] = {V1[3 * i], V1[3 * i + 1], V1[3 * i + 2]};
  const long long ip = (i == 0) ? 0 : i - 1;   // np.roll(.,1) with [0] = itself (simple.py:36-39)
  T x[18];
  cat_features18<T, M>(p0, p1, v1, a1, theta[i], gamma[i], theta[ip], gamma[ip], x);
#pragma unroll
  for (int j = 0; j < 18; ++j)
    X[i * 18 + j] = apply_scaler ? (x[j] - CatScaler<T>::mean(j)) * CatScaler<T>::inv_scale(j) : x[j];
}

// ------------------------------------------------------------------------------------------
// Feature term bank: every per-row term the reference's other feature builders stack
// -- build_theta_features / build_gamma_features (main_fun.py:267-415: cable-relative surge/sway, span l,
// elevation dH, catenary tension T = (w/L) l / (2 sinh(C l / 2)) with C from solve_catenary and the
// nan -> (w/L) l / 2 fallback, tanh of the bounded projection, lateral acceleration), build_clean_features
// (main_fun.py:197-265) and extract_features (main_fun.py:166-193) -- computed once per row, including the
// per-row catenary solve the reference runs through np.vectorize(brentq).  The host wrappers assemble each
// layout by column selection.  A1 = np.gradient(V1, time) as in k_features.
// Columns: 0-2 P1 | 3-5 V1 | 6-8 A1 | 9-11 unit_rel | 12 clip(|rel|,1e-5,10) | 13 angle_proj | 14 v_surge
//   | 15 v_sway | 16 l | 17 dH | 18 C | 19 T | 20 tanh(angle_proj) | 21 A1.unit_rel | 22 |A1_perp| | 23 |V1|
//   | 24 |A1| | 25 |A1|/(|V1|+1e-8) | 26 theta | 27 gamma | 28 theta_prev | 29 gamma_prev | 30 |rel| | 31 0
// ------------------------------------------------------------------------------------------
#define CAT_BANK_W 32
template <typename T>
__global__ void __launch_bounds__(128) k_feature_bank(long long T_len, const T* __restrict__ P0,
                                                      const T* __restrict__ P1, const T* __restrict__ V1,
                                                      const T* __restrict__ time, const T* __restrict__ theta,
                                                      const T* __restrict__ gamma, T L, T w, T c_lo, T c_hi,
                                                      T* __restrict__ bank) {
  typedef CatMath<T> M;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= T_len) return;
  T a1[3];
  if (T_len < 2) {
    a1[0] = a1[1] = a1[2] = T(0);
  } else if (i == 0) {
    const T h = time[1] - time[0];
#pragma unroll
    for (int j = 0; j < 3; ++j) a1[j] = (V1[3 + j] - V1[j]) / h;
  } else if (i == T_len - 1) {
    const T h = time[i] - time[i - 1];
#pragma unroll
    for (int j = 0; j < 3; ++j) a1[j] = (V1[3 * i + j] - V1[3 * (i - 1) + j]) / h;
  } else {
    const T hs = time[i] - time[i - 1], hd = time[i + 1] - time[i];
    const T a = -hd / (hs * (hd + hs)), b = (hd - hs) / (hd * hs), c = hs / (hd * (hd + hs));
#pragma unroll
    for (int j = 0; j < 3; ++j) a1[j] = a * V1[3 * (i - 1) + j] + b * V1[3 * i + j] + c * V1[3 * (i + 1) + j];
  }
  const T p1[3] = {P1[3 * i], P1[3 * i + 1], P1[3 * i + 2]};
  const T v[3] = {V1[3 * i], V1[3 * i + 1], V1[3 * i + 2]};
  const T rx = p1[0] - P0[3 * i], ry = p1[1] - P0[3 * i + 1], rz = p1[2] - P0[3 * i + 2];
  const T l2 = rx * rx + ry * ry;
  const T nrm = sqrt(l2 + rz * rz);
  const T den = nrm + T(1e-8);
  const T u[3] = {rx / den, ry / den, rz / den};
  const T dot = v[0] * u[0] + v[1] * u[1] + v[2] * u[2];
  const T vn = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  T ap = dot / (vn + T(1e-8));
  ap = ap < T(-1) ? T(-1) : (ap > T(1) ? T(1) : ap);
  const T cx = v[1] * u[2] - v[2] * u[1], cy = v[2] * u[0] - v[0] * u[2], cz = v[0] * u[1] - v[1] * u[0];
  const T l = sqrt(l2);
  const CatSpan<T> sp = cat_span<T, M>(l, rz, L, c_lo, c_hi, cat_cold<T>());
  const T wl = w / L;
  const T tens = (sp.C == sp.C) ? (wl * l) / (T(2) * sp.sh) : wl * l / T(2);
  const T adot = a1[0] * u[0] + a1[1] * u[1] + a1[2] * u[2];
  const T px = a1[0] - adot * u[0], py = a1[1] - adot * u[1], pz = a1[2] - adot * u[2];
  const T an = sqrt(a1[0] * a1[0] + a1[1] * a1[1] + a1[2] * a1[2]);
  const long long ip = (i == 0) ? 0 : i - 1;
  T* o = bank + i * CAT_BANK_W;
  o[0] = p1[0]; o[1] = p1[1]; o[2] = p1[2];
  o[3] = v[0]; o[4] = v[1]; o[5] = v[2];
  o[6] = a1[0]; o[7] = a1[1]; o[8] = a1[2];
  o[9] = u[0]; o[10] = u[1]; o[11] = u[2];
  o[12] = nrm < T(1e-5) ? T(1e-5) : (nrm > T(10) ? T(10) : nrm);
  o[13] = ap;
  o[14] = dot;
  o[15] = sqrt(cx * cx + cy * cy + cz * cz);
  o[16] = l;
  o[17] = rz;
  o[18] = sp.C;
  o[19] = tens;
  o[20] = tanh(ap);
  o[21] = adot;
  o[22] = sqrt(px * px + py * py + pz * pz);
  o[23] = vn;
  o[24] = an;
  o[25] = an / (vn + T(1e-8));
  o[26] = theta[i];
  o[27] = gamma[i];
  o[28] = theta[ip];
  o[29] = gamma[ip];
  o[30] = nrm;
  o[31] = T(0);
}

// uniform_resample (main_fun.py:128-145) = scipy interp1d(kind='linear', fill_value='extrapolate') of every
// column onto a new time grid.  One thread per output row: idx = lower_bound(time_in, t) clipped to
// [1, n_in-1] (np.searchsorted side='left' as scipy does), then scipy's (1.18) symmetric form
// ((t - x_lo)/(x_hi - x_lo)) y_hi + ((x_hi - t)/(x_hi - x_lo)) y_lo with every product and the sum rounded
// separately (no FMA contraction), so the result is bit-identical to interp1d.
// cols / out are row-major (rows, n_cols) like DataFrame.values.
template <typename T>
__global__ void __launch_bounds__(256) k_resample_linear(long long n_in, const T* __restrict__ time_in, int n_cols,
                                                         const T* __restrict__ cols, long long n_out,
                                                         const T* __restrict__ time_out, T* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_out) return;
  const T t = time_out[i];
  long long lo = 0, hi = n_in;   // first index with time_in[idx] >= t
  while (lo < hi) {
    const long long mid = (lo + hi) >> 1;
    if (time_in[mid] < t) lo = mid + 1;
    else hi = mid;
  }
  long long idx = lo < 1 ? 1 : (lo > n_in - 1 ? n_in - 1 : lo);
  const T x_lo = time_in[idx - 1], x_hi = time_in[idx];
  const T dx = x_hi - x_lo;
  const T w_hi = (t - x_lo) / dx, w_lo = (x_hi - t) / dx;
  const T* y_lo = cols + (idx - 1) * n_cols;
  const T* y_hi = cols + idx * n_cols;
  T* o = out + i * n_cols;
  for (int c = 0; c < n_cols; ++c) o[c] = __dadd_rn(__dmul_rn(w_hi, y_hi[c]), __dmul_rn(w_lo, y_lo[c]));
}
