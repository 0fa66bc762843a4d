// cat_kernels_expr.cuh -- run-time equation evaluator (sm_100a).
//
// The build-time code generator (codegen.py -> gen/dyn_eqs.cuh) covers the equations shipped in
// saved_models/.  The reference also holds ~3,900 hall-of-fame rows from its other PySR runs
// (cluster_run/outputs/*/d*_results_*.csv, outputs/*/hall_of_fame_*.csv) over other feature layouts, and
// a user retrains at will (simple.py:60-75).  Those go through this evaluator without rebuilding the
// library: the host compiles the equation string (the CSV's sympy_format / Equation column -- what
// PySRRegressor.predict lambdifies) into a postfix program; each thread runs it on its feature row with a
// small private stack.  Operator vocabulary = everything that occurs in the reference's equation files:
// + - * / neg sin cos tanh exp log sqrt abs square cube sign, integer and real powers.
#pragma once
#include "cat_math.cuh"

#define CAT_EXPR_MAX_OPS 2048
#define CAT_EXPR_STACK 32

enum {
  CAT_OP_CONST = 0,   // push consts[arg]
  CAT_OP_VAR = 1,     // push x[arg]
  CAT_OP_ADD = 2,
  CAT_OP_SUB = 3,
  CAT_OP_MUL = 4,
  CAT_OP_DIV = 5,
  CAT_OP_NEG = 6,
  CAT_OP_SIN = 7,
  CAT_OP_COS = 8,
  CAT_OP_TANH = 9,
  CAT_OP_EXP = 10,
  CAT_OP_LOG = 11,
  CAT_OP_SQRT = 12,
  CAT_OP_ABS = 13,
  CAT_OP_SQUARE = 14,
  CAT_OP_CUBE = 15,
  CAT_OP_SIGN = 16,
  CAT_OP_POW = 17,    // a ** b, both on the stack (numpy float power semantics)
  CAT_OP_POWI = 18,   // a ** arg, integer arg (repeated multiplication like sympy's lambdify -> numpy)
  CAT_OP_COUNT = 19
};

template <typename T> struct CatExprFn;
template <> struct CatExprFn<double> {
  static __device__ __forceinline__ double sin_(double x) { return sin(x); }
  static __device__ __forceinline__ double cos_(double x) { return cos(x); }
  static __device__ __forceinline__ double tanh_(double x) { return tanh(x); }
  static __device__ __forceinline__ double exp_(double x) { return exp(x); }
  static __device__ __forceinline__ double log_(double x) { return log(x); }
  static __device__ __forceinline__ double sqrt_(double x) { return sqrt(x); }
  static __device__ __forceinline__ double pow_(double a, double b) { return pow(a, b); }
};
template <> struct CatExprFn<float> {
  static __device__ __forceinline__ float sin_(float x) { return sinf(x); }
  static __device__ __forceinline__ float cos_(float x) { return cosf(x); }
  static __device__ __forceinline__ float tanh_(float x) { return tanhf(x); }
  static __device__ __forceinline__ float exp_(float x) { return expf(x); }
  static __device__ __forceinline__ float log_(float x) { return logf(x); }
  static __device__ __forceinline__ float sqrt_(float x) { return sqrtf(x); }
  static __device__ __forceinline__ float pow_(float a, float b) { return powf(a, b); }
};

// Feature access of one evaluation: x[j] = a[j]                 (b == nullptr)
//                                          (a[j] + b[j]) / 2     (the RK4 midpoint row of
//                                                                 simulate_rk4_theta_gamma.py:62)
// optionally standardised: (x - mean[j]) * inv_scale[j]  (test_temp.py:49)
template <typename T> struct CatExprRow {
  const T* a;
  const T* b;
  const T* mean;
  const T* inv_scale;
  __device__ __forceinline__ T get(int j) const {
    T v = a[j];
    if (b) v = (v + b[j]) / T(2);
    if (mean) v = (v - mean[j]) * inv_scale[j];
    return v;
  }
};

// ops: (opcode, arg) pairs in shared memory; consts in shared memory
template <typename T>
__device__ __forceinline__ T cat_expr_run(const int2* ops, int n_ops, const T* consts, const CatExprRow<T>& row) {
  typedef CatExprFn<T> F;
  T st[CAT_EXPR_STACK];
  int sp = 0;   // number of values on the stack; the host validated 1 <= depth <= CAT_EXPR_STACK throughout
  for (int i = 0; i < n_ops; ++i) {
    const int2 o = ops[i];
    switch (o.x) {
      case CAT_OP_CONST: st[sp++] = consts[o.y]; break;
      case CAT_OP_VAR: st[sp++] = row.get(o.y); break;
      case CAT_OP_ADD: --sp; st[sp - 1] = st[sp - 1] + st[sp]; break;
      case CAT_OP_SUB: --sp; st[sp - 1] = st[sp - 1] - st[sp]; break;
      case CAT_OP_MUL: --sp; st[sp - 1] = st[sp - 1] * st[sp]; break;
      case CAT_OP_DIV: --sp; st[sp - 1] = st[sp - 1] / st[sp]; break;
      case CAT_OP_NEG: st[sp - 1] = -st[sp - 1]; break;
      case CAT_OP_SIN: st[sp - 1] = F::sin_(st[sp - 1]); break;
      case CAT_OP_COS: st[sp - 1] = F::cos_(st[sp - 1]); break;
      case CAT_OP_TANH: st[sp - 1] = F::tanh_(st[sp - 1]); break;
      case CAT_OP_EXP: st[sp - 1] = F::exp_(st[sp - 1]); break;
      case CAT_OP_LOG: st[sp - 1] = F::log_(st[sp - 1]); break;
      case CAT_OP_SQRT: st[sp - 1] = F::sqrt_(st[sp - 1]); break;
      case CAT_OP_ABS: st[sp - 1] = st[sp - 1] < T(0) ? -st[sp - 1] : (st[sp - 1] == T(0) ? T(0) : st[sp - 1]); break;
      case CAT_OP_SQUARE: st[sp - 1] = st[sp - 1] * st[sp - 1]; break;
      case CAT_OP_CUBE: st[sp - 1] = st[sp - 1] * st[sp - 1] * st[sp - 1]; break;
      case CAT_OP_SIGN: {
        const T v = st[sp - 1];
        st[sp - 1] = v > T(0) ? T(1) : (v < T(0) ? T(-1) : v);   // 0 -> 0, NaN -> NaN (np.sign)
        break;
      }
      case CAT_OP_POW: --sp; st[sp - 1] = F::pow_(st[sp - 1], st[sp]); break;
      case CAT_OP_POWI: {
        const T b = st[sp - 1];
        int n = o.y < 0 ? -o.y : o.y;
        T r = T(1), m = b;
        while (n) {   // binary powering; numpy's integer-exponent pow is pow(): equal to ~1 ulp
          if (n & 1) r = r * m;
          m = m * m;
          n >>= 1;
        }
        st[sp - 1] = o.y < 0 ? T(1) / r : r;
        break;
      }
      default: break;
    }
  }
  return st[0];
}

// block-cooperative staging of the program into shared memory
template <typename T>
__device__ __forceinline__ void cat_expr_stage(const int2* __restrict__ ops, int n_ops, const double* __restrict__ consts,
                                               int n_consts, int2* s_ops, T* s_consts) {
  for (int i = threadIdx.x; i < n_ops; i += blockDim.x) s_ops[i] = ops[i];
  for (int i = threadIdx.x; i < n_consts; i += blockDim.x) s_consts[i] = T(consts[i]);
  __syncthreads();
}

// predict: out[i] = f(X[i, :F])
template <typename T>
__global__ void __launch_bounds__(256) k_expr_eval(long long n, int F, const T* __restrict__ X,
                                                   const int2* __restrict__ ops, int n_ops,
                                                   const double* __restrict__ consts, int n_consts,
                                                   const T* __restrict__ mean, const T* __restrict__ inv_scale,
                                                   T* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char cat_expr_smem[];
  int2* s_ops = reinterpret_cast<int2*>(cat_expr_smem);
  T* s_consts = reinterpret_cast<T*>(s_ops + n_ops);
  cat_expr_stage<T>(ops, n_ops, consts, n_consts, s_ops, s_consts);
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  CatExprRow<T> row;
  row.a = X + i * F;
  row.b = nullptr;
  row.mean = mean;
  row.inv_scale = inv_scale;
  out[i] = cat_expr_run<T>(s_ops, n_ops, s_consts, row);
}

// open-loop step increments, as k_rk4_increments but for a run-time program and any feature count:
// method 0 = RK4 (simulate_rk4_theta_gamma.py:55-66), 1 = Euler (main_fun.py:758-762),
// 2 = the slope itself times dt, f(X[i-1]) * dt (first pass of the second-order scheme, test_cluster.py:114-117)
template <typename T>
__global__ void __launch_bounds__(256) k_expr_increments(long long n_traj, long long T_len, int F,
                                                         const T* __restrict__ X, const T* __restrict__ time,
                                                         const int2* __restrict__ ops, int n_ops,
                                                         const double* __restrict__ consts, int n_consts,
                                                         int method, T* __restrict__ inc) {
  extern __shared__ __align__(16) unsigned char cat_expr_smem[];
  int2* s_ops = reinterpret_cast<int2*>(cat_expr_smem);
  T* s_consts = reinterpret_cast<T*>(s_ops + n_ops);
  cat_expr_stage<T>(ops, n_ops, consts, n_consts, s_ops, s_consts);
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long per = T_len - 1;
  if (g >= n_traj * per) return;
  const long long tr = g / per;
  const long long i = g - tr * per + 1;
  const T dt = time[i] - time[i - 1];
  CatExprRow<T> row;
  row.a = X + (tr * T_len + i - 1) * F;
  row.b = nullptr;
  row.mean = nullptr;
  row.inv_scale = nullptr;
  const T k1 = cat_expr_run<T>(s_ops, n_ops, s_consts, row);
  T r;
  if (method != 0) {
    r = k1 * dt;
  } else {
    row.b = X + (tr * T_len + i) * F;
    const T k2 = cat_expr_run<T>(s_ops, n_ops, s_consts, row);
    const T k3 = k2;
    row.a = row.b;
    row.b = nullptr;
    const T k4 = cat_expr_run<T>(s_ops, n_ops, s_consts, row);
    r = (dt / T(6)) * (k1 + T(2) * k2 + T(2) * k3 + k4);
  }
  inc[tr * T_len + i] = r;
}

// second pass of the second-order scheme (test_cluster.py:125-128): inc2[i] = ydot[i-1] * dt
template <typename T>
__global__ void __launch_bounds__(256) k_expr_second_pass(long long n_traj, long long T_len,
                                                          const T* __restrict__ time, const T* __restrict__ ydot,
                                                          T* __restrict__ inc) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long per = T_len - 1;
  if (g >= n_traj * per) return;
  const long long tr = g / per;
  const long long i = g - tr * per + 1;
  inc[tr * T_len + i] = ydot[tr * T_len + i - 1] * (time[i] - time[i - 1]);
}
