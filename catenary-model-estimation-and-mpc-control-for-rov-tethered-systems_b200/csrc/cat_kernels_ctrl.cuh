// cat_kernels_ctrl.cuh -- receding-horizon controller around k_mpc_rollout (sm_100a).
//
// One control step = k_mpc_rollout (candidates sampled around the nominal sequence, argmin)
//                    [+ k_ctrl_mppi: exp(-(J-Jmin)/lambda)-weighted mean of all candidates]
//                    + k_ctrl_advance: next nominal sequence (shifted by one step), plant step with its
//                      first control, log row, step counter.
// Plant state, nominal sequence, step counter and log never leave HBM, so the step is replayed as one CUDA
// graph.  The reference has no controller (its pympc submodule is empty); the definition is
// oracle/catenary_oracle.py::mpc_receding_horizon and these kernels follow it.
#pragma once
#include "cat_kernels_mpc.cuh"

struct CatCtrlDev {
  int mode;        // CAT_CTRL_*
  int max_steps;   // log capacity
  int slices;      // MPPI: candidate slices per horizon step (grid.z of k_ctrl_mppi), >= 1
  double lambda;
};

// deterministic block sum of 4 doubles (fixed shuffle tree, then warp order); result in thread 0
__device__ __forceinline__ void cat_block_sum4(double (&v)[4]) {
  __shared__ double s_part[8][4];
#pragma unroll
  for (int off = 16; off > 0; off >>= 1)
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] += __shfl_xor_sync(0xffffffffu, v[j], off);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0)
#pragma unroll
    for (int j = 0; j < 4; ++j) s_part[w][j] = v[j];
  __syncthreads();
  if (threadIdx.x == 0) {
    const int nw = blockDim.x >> 5;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      double a = s_part[0][j];
      for (int i = 1; i < nw; ++i) a += s_part[i][j];
      v[j] = a;
    }
  }
}

// MPPI blend: u_star[prob][h][:] = sum_k w_k u_k[h] / sum_k w_k, w_k = exp(-(J_k - Jmin)/lambda), NaN J -> 0.
// grid (H, n_problems, slices), 256 threads: block (h, prob, z) sums slice z of the candidates into
// partial[prob][h][z][4] = (sum w u_x, sum w u_y, sum w u_z, sum w); k_ctrl_advance adds the slices in order.
// The candidates' controls are regenerated from the Philox counter, not stored.
template <typename T>
__global__ void __launch_bounds__(256) k_ctrl_mppi(const CatMpcDev<T> q, const CatCtrlDev x,
                                                   const T* __restrict__ costs,
                                                   const double* __restrict__ best_cost,
                                                   const long long* __restrict__ best_idx,
                                                   double* __restrict__ partial) {
  typedef CatMath<T> M;
  __shared__ CatIcdfSmem<T> s_icdf_tab;
  cat_icdf_stage<T>(s_icdf_tab);
  const CatIcdfRef s_icdf = cat_icdf_ref_of<T>(s_icdf_tab);
  const int h = blockIdx.x, prob = blockIdx.y, z = blockIdx.z;
  const unsigned long long seed_eff = q.seed + (unsigned long long)(*q.step);
  const double jmin = best_cost[prob];
  double* out = partial + (((long long)prob * q.H + h) * x.slices + z) * 4;
  if (!(jmin == jmin)) {   // NaN winner: fall back to the best candidate itself (weight 1 in slice 0)
    if (threadIdx.x == 0) {
      T u[3] = {T(0), T(0), T(0)};
      if (z == 0) cat_nominal_control<T, M>(q, seed_eff, prob, best_idx[prob], h, s_icdf, q.clip, u);
      out[0] = (double)u[0];
      out[1] = (double)u[1];
      out[2] = (double)u[2];
      out[3] = z == 0 ? 1.0 : 0.0;
    }
    return;
  }
  const long long per = ((q.K + x.slices - 1) / x.slices + 255) / 256 * 256;
  const long long k_lo = (long long)z * per;
  const long long k_hi = k_lo + per < q.K ? k_lo + per : q.K;
  const double inv_l = -1.0 / x.lambda;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  const T* jc = costs + (long long)prob * q.K;
  for (long long k = k_lo + threadIdx.x; k < k_hi; k += blockDim.x) {
    const double j = (double)jc[k];
    double w = exp((j - jmin) * inv_l);
    if (!(j == j)) w = 0.0;
    T u[3];
    cat_nominal_control<T, M>(q, seed_eff, prob, q.k_offset + k, h, s_icdf, q.clip, u);
    acc[0] += w * (double)u[0];
    acc[1] += w * (double)u[1];
    acc[2] += w * (double)u[2];
    acc[3] += w;
  }
  cat_block_sum4(acc);
  if (threadIdx.x == 0) {
    out[0] = acc[0];
    out[1] = acc[1];
    out[2] = acc[2];
    out[3] = acc[3];
  }
}

// one fp64 step of the plant (= the rollout model) with the applied control u; also the step's cost terms.
// Same operations, in the same association, as one horizon step of oracle mpc_rollout_costs.
__device__ __forceinline__ void cat_plant_step(const CatMpcDev<double>& q, CatMpcProblem& s, const double u[3],
                                               const double* __restrict__ ref, double* __restrict__ row) {
  typedef CatMath<double> M;
  const double P0[3] = {s.P0[0], s.P0[1], s.P0[2]};
  double Pn[3], An[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    Pn[a] = s.P1[a] + u[a] * q.dt_mm;
    An[a] = (u[a] - s.V1[a]) / q.dt;
  }
  double x0[18], x1[18], xm[18];
  cat_features18<double, M>(P0, s.P1, s.V1, s.A1, s.theta, s.gamma, s.theta_prev, s.gamma_prev, x0);
  cat_scale18<double>(x0);
  const double k1t = cat_eval_equation<double, M>(q.eq_theta, x0);
  const double k1g = cat_eval_equation<double, M>(q.eq_gamma, x0);
  cat_features18<double, M>(P0, Pn, u, An, s.theta + q.dt * k1t, s.gamma + q.dt * k1g, s.theta, s.gamma, x1);
  cat_scale18<double>(x1);
#pragma unroll
  for (int j = 0; j < 18; ++j) xm[j] = (x0[j] + x1[j]) * 0.5;
  const double k2t = cat_eval_equation<double, M>(q.eq_theta, xm);
  const double k2g = cat_eval_equation<double, M>(q.eq_gamma, xm);
  const double k4t = cat_eval_equation<double, M>(q.eq_theta, x1);
  const double k4g = cat_eval_equation<double, M>(q.eq_gamma, x1);
  const double thn = s.theta + q.dt6 * (k1t + 2.0 * k2t + 2.0 * k2t + k4t);
  const double gan = s.gamma + q.dt6 * (k1g + 2.0 * k2g + 2.0 * k2g + k4g);
  const double rx = Pn[0] - P0[0], ry = Pn[1] - P0[1], rz = Pn[2] - P0[2];
  const double l2 = rx * rx + ry * ry;
  const double l = M::sqrt_(l2);
  const double d = M::sqrt_(l2 + rz * rz);
  const CatSpan<double> sp = cat_span<double, M, true>(l, rz, q.L, q.c_lo, q.c_hi, cat_cold<double>());
  const double z_low = cat_z_low<double, M>(sp, l, q.L, P0[2], P0[2] + rz);
  const double ex = Pn[0] - ref[0], ey = Pn[1] - ref[1], ez = Pn[2] - ref[2];
  const double dvx = u[0] - s.V1[0], dvy = u[1] - s.V1[1], dvz = u[2] - s.V1[2];
  const double over = M::pos_(d - q.rhoL);
  const double deep = M::pos_(q.z_thr - z_low);
  double c = q.w_trk * (ex * ex + ey * ey + ez * ez);
  c = c + q.w_du * (dvx * dvx + dvy * dvy + dvz * dvz);
  c = c + q.w_len * (over * over);
  c = c + q.w_sag * (deep * deep);
  c = c + q.w_ang * (thn * thn + gan * gan);
  if (!(sp.C == sp.C)) c = c + q.pen;
  s.theta_prev = s.theta;
  s.gamma_prev = s.gamma;
  s.theta = thn;
  s.gamma = gan;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    s.P1[a] = Pn[a];
    s.V1[a] = u[a];
    s.A1[a] = An[a];
  }
  row[0] = Pn[0];
  row[1] = Pn[1];
  row[2] = Pn[2];
  row[3] = u[0];
  row[4] = u[1];
  row[5] = u[2];
  row[6] = thn;
  row[7] = gan;
  row[8] = sp.C;
  row[9] = z_low;
  row[12] = c;
  row[14] = d;
  row[15] = 0.0;
}

// grid (n_problems), 64 threads, dynamic shared memory H*3 doubles.
// qd carries the fp64 plant constants; q the rollout-type sampler constants (same values, type T).
template <typename T>
__global__ void __launch_bounds__(64) k_ctrl_advance(const CatMpcDev<T> q, const CatMpcDev<double> qd,
                                                     const CatCtrlDev x, CatMpcProblem* __restrict__ problems,
                                                     double* __restrict__ u_nom,
                                                     const double* __restrict__ u_star,
                                                     const double* __restrict__ p_ref,
                                                     const double* __restrict__ best_cost,
                                                     const long long* __restrict__ best_idx,
                                                     double* __restrict__ log, int* __restrict__ step,
                                                     unsigned int* __restrict__ ticket) {
  typedef CatMath<T> M;
  extern __shared__ double s_u[];   // [H][3]
  __shared__ CatIcdfSmem<T> s_icdf_tab;
  if (x.mode != CAT_CTRL_MPPI) cat_icdf_stage<T>(s_icdf_tab);   // block-uniform
  const CatIcdfRef s_icdf = cat_icdf_ref_of<T>(s_icdf_tab);
  const int prob = blockIdx.x;
  const int t = *step;
  const unsigned long long seed_eff = q.seed + (unsigned long long)t;
  const long long bi = best_idx[prob];
  for (int h = threadIdx.x; h < q.H; h += blockDim.x) {
    if (x.mode == CAT_CTRL_MPPI) {
      const double* ps = u_star + ((long long)prob * q.H + h) * x.slices * 4;   // slice partial sums
      double a0 = ps[0], a1 = ps[1], a2 = ps[2], aw = ps[3];
      for (int z = 1; z < x.slices; ++z) {
        a0 += ps[4 * z + 0];
        a1 += ps[4 * z + 1];
        a2 += ps[4 * z + 2];
        aw += ps[4 * z + 3];
      }
      s_u[3 * h + 0] = a0 / aw;
      s_u[3 * h + 1] = a1 / aw;
      s_u[3 * h + 2] = a2 / aw;
    } else {
      T u[3];
      cat_nominal_control<T, M>(q, seed_eff, prob, bi, h, s_icdf, q.clip, u);
      s_u[3 * h + 0] = (double)u[0];
      s_u[3 * h + 1] = (double)u[1];
      s_u[3 * h + 2] = (double)u[2];
    }
  }
  __syncthreads();   // every read of the old nominal sequence is done
  double* un = u_nom + (long long)prob * q.H * 3;
  for (int h = threadIdx.x; h < q.H; h += blockDim.x) {
    const int src = h + 1 < q.H ? h + 1 : q.H - 1;
    un[3 * h + 0] = s_u[3 * src + 0];
    un[3 * h + 1] = s_u[3 * src + 1];
    un[3 * h + 2] = s_u[3 * src + 2];
  }
  if (threadIdx.x == 0) {
    CatMpcProblem s = problems[prob];
    const double u0[3] = {s_u[0], s_u[1], s_u[2]};
    const int tr = t < qd.ref_len ? t : qd.ref_len - 1;
    const double* ref = p_ref + ((long long)prob * qd.ref_len + tr) * 3;
    double row[CAT_CTRL_LOG_W];
    cat_plant_step(qd, s, u0, ref, row);
    row[10] = best_cost[prob];
    row[11] = (double)bi;
    row[13] = (double)t;
    problems[prob] = s;
    if (t < x.max_steps) {
      double* lr = log + ((long long)prob * x.max_steps + t) * CAT_CTRL_LOG_W;
#pragma unroll
      for (int j = 0; j < CAT_CTRL_LOG_W; ++j) lr[j] = row[j];
    }
    // the last block to get here (all blocks have read *step by now) advances the counter
    __threadfence();
    if (atomicAdd(ticket, 1u) == gridDim.x - 1) {
      *ticket = 0u;
      *step = t + 1;
    }
  }
}
