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
