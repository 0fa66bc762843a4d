// catenary_b200.cu -- C ABI of libcatenary_b200.so (include/catenary_b200.h), sm_100a only.
//
// Host side of the boundary: context (device, stream, scratch), host<->device staging for the
// host-pointer entry points, kernel launches.  No CPU fallback anywhere: every compute entry point
// launches a CUDA kernel or fails with CAT_ECUDA.
#include <cuda_runtime.h>

#include <mutex>
#include <nvtx3/nvToolsExt.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "cat_host.cuh"
#include "cat_kernels_dyn.cuh"
#include "cat_kernels_geom.cuh"
#include "cat_mpc_launch.cuh"
#include "cat_kernels_ctrl.cuh"
#include "cat_kernels_expr.cuh"
#include "cat_kernels_feat.cuh"

void cat_nvtx_push(const CatCtx*, const char* name) { nvtxRangePushA(name); }
void cat_nvtx_pop(const CatCtx*) { nvtxRangePop(); }

namespace {

int stage(CatCtx* c) {
  if (c->h_stage) return CAT_OK;
  void* h = nullptr;
  cudaError_t e = cudaHostAlloc(&h, kStageBytes, cudaHostAllocMapped | cudaHostAllocPortable);
  if (e != cudaSuccess) return fail(c, CAT_ENOMEM, "cudaHostAlloc", e);
  void* d = nullptr;
  e = cudaHostGetDevicePointer(&d, h, 0);
  if (e != cudaSuccess) {
    cudaFreeHost(h);
    return fail(c, CAT_ECUDA, "cudaHostGetDevicePointer", e);
  }
  c->h_stage = (unsigned char*)h;
  c->d_stage = (unsigned char*)d;
  return CAT_OK;
}

// private per-block winner slots + tickets of one in-flight MPC step (cat_mpc_workspace_create)
struct CatMpcWs {
  CatBest* bb;
  unsigned int* tk;
  size_t n_bb, n_tk;
};

const double kCLo = 1e-6, kCHi = 10.0;   // bracket of main_fun.py:425

// ---------------------------------------------------------------------------------------------
template <typename T>
int solve_param_dev(CatCtx* c, int64_t n, const T* l, const T* dH, const T* L, int64_t L_stride, T* C_out) {
  if (n == 0) return CAT_OK;
  k_solve_param<T><<<nblk(n, 256), 256, 0, c->stream>>>(n, l, dH, L, L_stride, T(kCLo), T(kCHi), C_out);
  LAUNCHED();
  return CAT_OK;
}

template <typename T>
int solve_param_host(CatCtx* c, int64_t n, const T* l, const T* dH, const T* L, int64_t L_stride, T* C_out) {
  NEED(n >= 0 && (n == 0 || (l && dH && L && C_out)) && (L_stride == 0 || L_stride == 1));
  if (n == 0) return CAT_OK;
  int rc = bind(c);
  if (rc) return rc;
  T *dl, *dh, *dL, *dC;
  const size_t nb = (size_t)n * sizeof(T);
  SCRATCH(SLOT_A, nb, dl);
  SCRATCH(SLOT_B, nb, dh);
  SCRATCH(SLOT_C, L_stride ? nb : sizeof(T), dL);
  SCRATCH(SLOT_D, nb, dC);
  CU(cudaMemcpyAsync(dl, l, nb, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dh, dH, nb, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dL, L, L_stride ? nb : sizeof(T), cudaMemcpyHostToDevice, c->stream));
  rc = solve_param_dev<T>(c, n, dl, dh, dL, L_stride, dC);
  if (rc) return rc;
  CU(cudaMemcpyAsync(C_out, dC, nb, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

inline int popcount4(int w) { return (w & 1) + ((w >> 1) & 1) + ((w >> 2) & 1) + ((w >> 3) & 1); }

// ---------------------------------------------------------------------------------------------
template <typename T>
int shape_dev_on(CatCtx* c, cudaStream_t st, int64_t n, const T* A, const T* B, const T* L, int64_t L_stride,
                 const T* theta, const T* gamma, int32_t n_pts, int32_t which, T* out, T* C_out, uint8_t* valid) {
  if (n == 0) return CAT_OK;
  const size_t smem = (size_t)CAT_SHAPE_TILE * cat_shape_record_len(which) * 2 * sizeof(T);
#define CAT_SHAPE_LAUNCH(W)                                                                                          \
  k_shape<T, W><<<nblk(n, CAT_SHAPE_TILE), CAT_SHAPE_TILE, smem, st>>>(n, A, B, L, L_stride, theta, gamma, n_pts,       \
                                                                      which, T(kCLo), T(kCHi), out, C_out, valid)
  // Single-curve masks are compiled in for fp64 (A/B on one B200, 1,048,576 x 100 points: final curve 0.525 ->
  // 0.448 ms, original curve 0.437 -> 0.423 ms).  The same specialisation is slower for all four curves (1.86 ->
  // 2.21 ms: the unrolled per-map code spills at the 96-register cap) and in fp32 (0.267 -> 0.278 ms, two
  // configurations per warp), so those take the run-time form.
  const int w = sizeof(T) == 8 ? which : 0;
  // Two stores per point (a 2-value vector + a scalar) instead of three scalars, in the run-time form, when at least
  // two of the rotated curves are asked for (they share a span, so their stores interleave): one B200, 1,048,576 x 100
  // points, all four curves 1.87 -> 1.57 ms in fp64 (0.99 of the HBM peak) and 1.34 -> 0.98 ms in fp32; masks 12 and 14
  // likewise.  With ONE curve per span it loses (final curve fp32 0.268 -> 0.309 ms, original + final fp64 0.885 ->
  // 0.991 ms), so those keep the scalar stores (profiles/r2_ab_shape_two_stores.txt).  Needs every curve on a
  // 2-value boundary: even n_pts, aligned `out`.  CAT_SHAPE_ST2=0/1 forces it off/on (A/B).
  static const int st2_env = [] { const char* e = std::getenv("CAT_SHAPE_ST2"); return e ? std::atoi(e) : -1; }();
  const bool st2_ok = (n_pts % 2 == 0) && (reinterpret_cast<uintptr_t>(out) % (2 * sizeof(T)) == 0);
  const bool st2 = st2_ok && (st2_env >= 0 ? st2_env != 0 : popcount4(which & 14) >= 2);
  switch (w) {
    case CAT_OUT_ORIGINAL: CAT_SHAPE_LAUNCH(1); break;     // Catenary.__call__ alone (catenary.py:25)
    case CAT_OUT_FINAL: CAT_SHAPE_LAUNCH(8); break;        // fully augmented curve (main_fun.py:111, [3])
    default:
      if (st2)
        k_shape<T, 0, true><<<nblk(n, CAT_SHAPE_TILE), CAT_SHAPE_TILE, smem, st>>>(
            n, A, B, L, L_stride, theta, gamma, n_pts, which, T(kCLo), T(kCHi), out, C_out, valid);
      else
        CAT_SHAPE_LAUNCH(0);
      break;
  }
#undef CAT_SHAPE_LAUNCH
  LAUNCHED();
  return CAT_OK;
}


template <typename T>
int shape_host(CatCtx* c, int64_t n, const T* A, const T* B, const T* L, int64_t L_stride, const T* theta,
               const T* gamma, int32_t n_pts, int32_t which, T* out, T* C_out, uint8_t* valid) {
  NEED(n >= 0 && n_pts >= 1 && which > 0 && which < 16 && (L_stride == 0 || L_stride == 1));
  NEED(n == 0 || (A && B && L && out));
  if (n == 0) return CAT_OK;
  int rc = bind(c);
  if (rc) return rc;
  const int nout = popcount4(which);
  CatNvtxRange range(c, "cat_shape (host): chunked H2D / k_shape / D2H");
  // chunked two-lane pipeline: lane j owns its own scratch; H2D, kernel, D2H are in-order per lane and the
  // two lanes overlap each other (copy engine of one with the SMs of the other).
  const int64_t chunk = 65536;
  const size_t blk_elems = (size_t)chunk * n_pts * 3;
  struct Lane {
    T *A, *B, *L, *th, *ga, *out, *C;
    uint8_t* v;
  } lane[2];
  const int slots[2][4] = {{SLOT_A, SLOT_B, SLOT_C, SLOT_D}, {SLOT_E, SLOT_F, SLOT_G, SLOT_H}};
  for (int j = 0; j < 2; ++j) {
    // one input arena per lane: A(3) B(3) L(1) th(1) ga(1) C(1) = 10 values per config, + valid bytes
    T* in;
    SCRATCH(slots[j][0], (size_t)chunk * 10 * sizeof(T), in);
    lane[j].A = in;
    lane[j].B = in + 3 * chunk;
    lane[j].L = in + 6 * chunk;
    lane[j].th = in + 7 * chunk;
    lane[j].ga = in + 8 * chunk;
    lane[j].C = in + 9 * chunk;
    SCRATCH(slots[j][1], (size_t)chunk, lane[j].v);
    const int64_t need = (n < chunk) ? n : chunk;
    SCRATCH(slots[j][2], (size_t)need * n_pts * 3 * sizeof(T) * nout, lane[j].out);
    if (n <= chunk) break;
  }
  (void)blk_elems;
  cudaStream_t st[2] = {c->stream, c->aux_stream};
  CU(cudaEventRecord(c->ev_aux, c->stream));
  CU(cudaStreamWaitEvent(c->aux_stream, c->ev_aux, 0));
  int ci = 0;
  for (int64_t base = 0; base < n; base += chunk, ++ci) {
    const int j = ci & 1;
    const int64_t m = (n - base < chunk) ? (n - base) : chunk;
    Lane& ln = lane[j];
    CU(cudaMemcpyAsync(ln.A, A + 3 * base, (size_t)m * 3 * sizeof(T), cudaMemcpyHostToDevice, st[j]));
    CU(cudaMemcpyAsync(ln.B, B + 3 * base, (size_t)m * 3 * sizeof(T), cudaMemcpyHostToDevice, st[j]));
    CU(cudaMemcpyAsync(ln.L, L + (L_stride ? base : 0), (L_stride ? (size_t)m : 1) * sizeof(T),
                       cudaMemcpyHostToDevice, st[j]));
    if (theta) CU(cudaMemcpyAsync(ln.th, theta + base, (size_t)m * sizeof(T), cudaMemcpyHostToDevice, st[j]));
    if (gamma) CU(cudaMemcpyAsync(ln.ga, gamma + base, (size_t)m * sizeof(T), cudaMemcpyHostToDevice, st[j]));
    rc = shape_dev_on<T>(c, st[j], m, ln.A, ln.B, ln.L, L_stride, theta ? ln.th : nullptr,
                         gamma ? ln.ga : nullptr, n_pts, which, ln.out, ln.C, ln.v);
    if (rc) return rc;
    const size_t blk_dev = (size_t)m * n_pts * 3;          // elements per output block of this chunk
    const size_t blk_host = (size_t)n * n_pts * 3;         // elements per output block of the caller
    for (int o = 0; o < nout; ++o)
      CU(cudaMemcpyAsync(out + o * blk_host + (size_t)base * n_pts * 3, ln.out + o * blk_dev,
                         blk_dev * sizeof(T), cudaMemcpyDeviceToHost, st[j]));
    if (C_out) CU(cudaMemcpyAsync(C_out + base, ln.C, (size_t)m * sizeof(T), cudaMemcpyDeviceToHost, st[j]));
    if (valid) CU(cudaMemcpyAsync(valid + base, ln.v, (size_t)m, cudaMemcpyDeviceToHost, st[j]));
  }
  CU(cudaStreamSynchronize(c->aux_stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

// ---------------------------------------------------------------------------------------------
template <typename T>
int predict_host(CatCtx* c, int64_t n, const T* X, int eq_id, int apply_scaler, T* out) {
  NEED(n >= 0 && eq_id >= 0 && eq_id < CAT_N_EQUATIONS && (n == 0 || (X && out)));
  if (n == 0) return CAT_OK;
  int rc = bind(c);
  if (rc) return rc;
  T *dX, *dO;
  SCRATCH(SLOT_A, (size_t)n * 18 * sizeof(T), dX);
  SCRATCH(SLOT_B, (size_t)n * sizeof(T), dO);
  CU(cudaMemcpyAsync(dX, X, (size_t)n * 18 * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  k_predict<T><<<nblk(n, 256), 256, 0, c->stream>>>(n, dX, eq_id, apply_scaler, dO);
  LAUNCHED();
  CU(cudaMemcpyAsync(out, dO, (size_t)n * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

template <typename T>
int rk4_open_host(CatCtx* c, int64_t n_traj, int64_t T_len, const T* X, const T* time, const T* y0, int eq_id,
                  int method, T* y_out) {
  NEED(n_traj >= 0 && T_len >= 0 && eq_id >= 0 && eq_id < CAT_N_EQUATIONS && (method == 0 || method == 1));
  if (n_traj == 0 || T_len == 0) return CAT_OK;
  NEED(X && time && y0 && y_out);
  int rc = bind(c);
  if (rc) return rc;
  T *dX, *dt, *dy0, *dinc, *dy;
  const size_t rows = (size_t)n_traj * T_len;
  SCRATCH(SLOT_A, rows * 18 * sizeof(T), dX);
  SCRATCH(SLOT_B, (size_t)T_len * sizeof(T), dt);
  SCRATCH(SLOT_C, (size_t)n_traj * sizeof(T), dy0);
  SCRATCH(SLOT_D, rows * sizeof(T), dinc);
  SCRATCH(SLOT_E, rows * sizeof(T), dy);
  CU(cudaMemcpyAsync(dX, X, rows * 18 * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dt, time, (size_t)T_len * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dy0, y0, (size_t)n_traj * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  if (T_len > 1) {
    const long long work = (long long)n_traj * (T_len - 1);
    k_rk4_increments<T><<<nblk(work, 256), 256, 0, c->stream>>>(n_traj, T_len, dX, dt, eq_id, method, dinc);
    LAUNCHED();
  }
  k_rk4_accumulate<T><<<nblk(n_traj, 128), 128, 0, c->stream>>>(n_traj, T_len, dy0, dinc, dy);
  LAUNCHED();
  CU(cudaMemcpyAsync(y_out, dy, rows * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

// ---------------------------------------------------------------------------------------------
// Threads per block of k_mpc_rollout: 128 = one warp per SM sub-partition per block.  Measured on B200 at
// config 3 (65,536 x 50, explicit controls): 128 -> 0.116 ms, 64 -> 0.120 ms, 32 -> 0.136 ms; smaller blocks
// balance the SMs better (512 blocks on 148 SMs is 3.46 per SM) but unbalance the four schedulers inside an SM,
// which costs more.  CAT_MPC_BLOCK_SIZE overrides (tuning only).
inline int mpc_block_size(CatCtx* c, long long K, int n_problems) {
  if (c->mpc_block) return c->mpc_block;
  (void)K;
  (void)n_problems;
  return 128;
}

template <typename T>
int mpc_fill(CatCtx* c, const CatMpcParams* p, CatMpcDev<T>& q) {
  NEED(p != nullptr);
  NEED(p->K >= 1 && p->H >= 1 && p->n_problems >= 1 && p->n_problems <= 65535);
  NEED(p->eq_theta >= -1 && p->eq_theta < CAT_N_EQUATIONS && p->eq_gamma >= -1 && p->eq_gamma < CAT_N_EQUATIONS);
  NEED(p->dt > 0 && p->L > 0);
  q.K = p->K;
  q.k_offset = p->k_offset;
  q.H = p->H;
  q.n_problems = p->n_problems;
  q.flags = p->flags;
  q.problem_offset = p->problem_offset;
  q.eq_theta = p->eq_theta < 0 ? CAT_EQ_DTHETA_SHIPPED : p->eq_theta;
  q.eq_gamma = p->eq_gamma < 0 ? CAT_EQ_DGAMMA_SHIPPED : p->eq_gamma;
  if (p->flags & CAT_MPC_CONTROLS_AOS) {   // [problem][K][H][3]
    q.sa = 1;
    q.sh = 3;
    q.sk = 3LL * p->H;
    q.sp = 3LL * p->H * p->K;
  } else {                                  // [problem][H][3][K]
    q.sk = 1;
    q.sa = p->K;
    q.sh = 3LL * p->K;
    q.sp = 3LL * p->H * p->K;
  }
  q.dt = T(p->dt);
  q.dt_mm = T(p->dt / 1000.0);
  q.dt6 = T(p->dt / 6);
  q.L = T(p->L);
  q.w_trk = T(p->w_trk);
  q.w_du = T(p->w_du);
  q.w_len = T(p->w_len);
  q.w_sag = T(p->w_sag);
  q.w_ang = T(p->w_ang);
  q.rhoL = T(p->rho * p->L);
  q.rhoL2 = q.rhoL > T(0) ? q.rhoL * q.rhoL : T(-1);   // a non-positive threshold is always 'active'
  q.L2 = q.L * q.L;
  q.carry_t = (CAT_EQ_TABLE[q.eq_theta].live_mask & 0xC000u) == 0;
  q.carry_g = (CAT_EQ_TABLE[q.eq_gamma].live_mask & 0xC000u) == 0;
  q.z_thr = T(p->z_floor + p->margin);
  q.c_lo = T(p->c_lo);
  q.c_hi = T(p->c_hi);
  q.pen = T(p->nan_penalty);
  q.clip = T(p->clip);
  q.deep_clip = q.clip > T(CAT_ICDF_SMEM_ZMAX);
  q.arrive = nullptr;
  for (int a = 0; a < 3; ++a) q.sigma[a] = T(p->sigma[a]);
  q.seed = p->seed;
  q.prog = nullptr;
  q.prog_consts = nullptr;
  q.n_ops_t = q.n_ops_g = q.n_c_t = q.n_c_g = 0;
  q.ref_smem_off = 0;
  q.p2p.peers = nullptr;
  q.p2p.rank = 0;
  q.p2p.n_ranks = 1;
  q.p2p.seq = 0;
  q.p2p.mode = CAT_P2P_LOCKSTEP;
  q.p2p.err = nullptr;
  q.u_nom = nullptr;
  q.step = nullptr;
  q.ref_len = p->H;
  for (int a = 0; a < 3; ++a) q.u_max[a] = T(0);
  return CAT_OK;
}

// bytes per stored control value of a step with these flags
template <typename T> inline size_t mpc_control_size(int flags) {
  return (sizeof(T) == 4 || (flags & CAT_MPC_CONTROLS_F32)) ? 4 : 8;
}

// launch of k_mpc_rollout for a filled q; bb/tk = per-block winner slots and tickets (the selected workspace,
// else the ctx scratch, when null)
template <typename T>
int mpc_launch(CatCtx* c, const CatMpcDev<T>& q, const CatMpcProblem* problems, const void* controls,
               const double* p_ref, double* best_cost, long long* best_idx, T* costs, CatBest* bb,
               unsigned int* tk) {
  // Launch shape: one thread per candidate.  (A role-split variant -- dynamics and catenary chains of a
  // candidate in two threads -- was measured in round 1 and was 13% SLOWER at config 3: the kernel is bound by
  // issue slots, not by the dependency chain, and the split duplicates loads and kinematics.  DESIGN.md 9.)
  CatMpcLaunchArgs a;
  a.block = mpc_block_size(c, q.K, q.n_problems);
  a.grid = dim3(nblk(q.K, a.block), (unsigned)q.n_problems, 1);
  a.stream = c->stream;
  a.rng = (q.flags & CAT_MPC_RNG_CONTROLS) != 0;
  a.controls_f32 = (q.flags & CAT_MPC_CONTROLS_F32) != 0;
  a.pdl = false;
  a.streamed = q.arrive != nullptr;
  if (c->overlap) {   // not inside a stream capture (the controller's graph): plain kernel nodes there
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(c->stream, &cs) == cudaSuccess && cs == cudaStreamCaptureStatusNone) a.pdl = true;
    else cudaGetLastError();
  }
  NEED(q.p2p.peers == nullptr || q.p2p.n_ranks <= a.block);
  const size_t n_bb = (size_t)a.grid.x * a.grid.y;
  if (!bb && c->ws) {
    CatMpcWs* w = (CatMpcWs*)c->ws;
    NEED(n_bb <= w->n_bb && (size_t)q.n_problems <= w->n_tk);
    bb = w->bb;
    tk = w->tk;
  }
  if (!bb) {
    void* pp = nullptr;
    int rc = scratch(c, SLOT_BEST, n_bb * sizeof(CatBest), &pp);
    if (rc) return rc;
    bb = (CatBest*)pp;
    rc = scratch(c, SLOT_TICKET, (size_t)65536 * sizeof(unsigned int), &pp, true);
    if (rc) return rc;
    tk = (unsigned int*)pp;
  }
  const bool shipped = (q.eq_theta == CAT_EQ_DTHETA_SHIPPED && q.eq_gamma == CAT_EQ_DGAMMA_SHIPPED);
  // dynamic shared memory: run-time programs (if any), then the H reference points in the compute type
  const size_t b_prog =
      q.prog ? (size_t)(q.n_ops_t + q.n_ops_g) * sizeof(int2) + (size_t)(q.n_c_t + q.n_c_g + 2) * sizeof(T) : 0;
  CatMpcDev<T> ql = q;
  ql.ref_smem_off = (int)((b_prog + 15) & ~(size_t)15);
  a.smem = (size_t)ql.ref_smem_off + (size_t)q.H * 3 * sizeof(T);
  if (a.smem > 48 * 1024)
    return fail(c, CAT_EINVAL, "horizon (and equation programs) do not fit the rollout's 48 KB of shared memory");
  int e;
  if (q.prog)   // run-time equation programs
    e = cat_mpc_launch_inst<T, CAT_KIND_VM>(a, ql, problems, controls, p_ref, costs, bb, tk, best_cost, best_idx);
  else if (shipped)
    e = cat_mpc_launch_inst<T, CAT_KIND_SHIPPED>(a, ql, problems, controls, p_ref, costs, bb, tk, best_cost, best_idx);
  else
    e = cat_mpc_launch_inst<T, CAT_KIND_GENERIC>(a, ql, problems, controls, p_ref, costs, bb, tk, best_cost, best_idx);
  ++c->launches;
  if (e != (int)cudaSuccess) return fail(c, CAT_ECUDA, "kernel launch (k_mpc_rollout)", (cudaError_t)e);
  return CAT_OK;
}

template <typename T>
int mpc_step_dev(CatCtx* c, const CatMpcParams* p, const CatMpcProblem* problems, const void* controls,
                 const double* p_ref, double* best_cost, long long* best_idx, T* costs) {
  CatMpcDev<T> q;
  int rc = mpc_fill<T>(c, p, q);
  if (rc) return rc;
  const bool rng = (p->flags & CAT_MPC_RNG_CONTROLS) != 0;
  NEED(problems && p_ref && best_cost && best_idx && (rng || controls));
  CatNvtxRange r(c, "cat_mpc_step_dev");
  return mpc_launch<T>(c, q, problems, controls, p_ref, best_cost, best_idx, costs, nullptr, nullptr);
}

// optional run-time equation programs of a host MPC step (both or neither)
struct CatHostPrograms {
  int32_t n_ops_t, n_c_t, n_ops_g, n_c_g;
  const int32_t *ops_t, *ops_g;
  const double *c_t, *c_g;
};

int expr_validate(CatCtx* c, int32_t n_features, int32_t n_ops, const int32_t* ops, int32_t n_consts,
                  const double* consts);

// upload [theta ops | gamma ops] and [theta consts | gamma consts] into SLOT_G and point q at them; the k1 carry
// holds for an equation that reads neither theta (column 14) nor gamma (column 15)
template <typename T>
int mpc_attach_programs(CatCtx* c, const CatHostPrograms* hp, CatMpcDev<T>& q) {
  int rc = expr_validate(c, 18, hp->n_ops_t, hp->ops_t, hp->n_c_t, hp->c_t);
  if (rc) return rc;
  rc = expr_validate(c, 18, hp->n_ops_g, hp->ops_g, hp->n_c_g, hp->c_g);
  if (rc) return rc;
  const size_t b_ops = (size_t)(hp->n_ops_t + hp->n_ops_g) * sizeof(int2);
  const size_t b_c = (size_t)(hp->n_c_t + hp->n_c_g + 1) * sizeof(double);
  // both programs are staged in the rollout's dynamic shared memory: keep the pair inside the 48 KB a kernel
  // gets without opting in (a pair of PySR equations is a few hundred bytes; expr_validate alone would admit 64 KB)
  if (b_ops + (size_t)(hp->n_c_t + hp->n_c_g + 2) * sizeof(T) > 48 * 1024)
    return fail(c, CAT_EINVAL, "equation programs too large for the rollout kernel (ops + constants > 48 KB)");
  unsigned char* base;
  SCRATCH(SLOT_G, b_ops + b_c, base);
  int2* d_ops = (int2*)base;
  double* d_c = (double*)(base + b_ops);
  CU(cudaMemcpyAsync(d_ops, hp->ops_t, (size_t)hp->n_ops_t * sizeof(int2), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d_ops + hp->n_ops_t, hp->ops_g, (size_t)hp->n_ops_g * sizeof(int2), cudaMemcpyHostToDevice,
                     c->stream));
  if (hp->n_c_t)
    CU(cudaMemcpyAsync(d_c, hp->c_t, (size_t)hp->n_c_t * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  if (hp->n_c_g)
    CU(cudaMemcpyAsync(d_c + hp->n_c_t, hp->c_g, (size_t)hp->n_c_g * sizeof(double), cudaMemcpyHostToDevice,
                       c->stream));
  q.prog = d_ops;
  q.prog_consts = d_c;
  q.n_ops_t = hp->n_ops_t;
  q.n_ops_g = hp->n_ops_g;
  q.n_c_t = hp->n_c_t;
  q.n_c_g = hp->n_c_g;
  auto reads_angles = [](int n, const int32_t* ops) {
    for (int i = 0; i < n; ++i)
      if (ops[2 * i] == CAT_OP_VAR && (ops[2 * i + 1] == 14 || ops[2 * i + 1] == 15)) return true;
    return false;
  };
  q.carry_t = reads_angles(hp->n_ops_t, hp->ops_t) ? 0 : 1;
  q.carry_g = reads_angles(hp->n_ops_g, hp->ops_g) ? 0 : 1;
  return CAT_OK;
}

// fused winner exchange over NVLink peer memory (CUDA IPC between the ranks of one node)
struct CatP2P {
  CatCtx* c;
  int rank, n_ranks;
  unsigned int seq;          // steps launched so far
  int mode;                  // CAT_P2P_MODE_*
  long long* slots;          // this rank's buffer [CAT_P2P_RING][n_ranks][CAT_P2P_SLOT_W]
  long long** d_peers;       // device array of n_ranks buffer addresses
  void* opened[CAT_P2P_MAX_RANKS];   // peers' mappings (null for self)
  int* d_err;
  bool connected;
};

inline void p2p_arm(const CatP2P* x, CatP2PDev& d) {   // the NEXT step's exchange parameters
  d.peers = x->d_peers;
  d.rank = x->rank;
  d.n_ranks = x->n_ranks;
  d.seq = x->seq + 1;
  d.mode = x->mode;
  d.err = x->d_err;
}

// Host-pointer MPC step.  x != null: the winners of the ranks are exchanged inside the step (lock-step) and
// best_cost / best_idx receive the JOB-WIDE winner (one problem).
template <typename T>
int mpc_step_host(CatCtx* c, const CatMpcParams* p, const CatMpcProblem* problems, const void* controls,
                  const double* p_ref, double* best_cost, int64_t* best_idx, T* costs,
                  const CatHostPrograms* hp = nullptr, CatP2P* x = nullptr) {
  NEED(p != nullptr && p->K >= 1 && p->H >= 1 && p->n_problems >= 1);
  const bool rng = (p->flags & CAT_MPC_RNG_CONTROLS) != 0;
  NEED(problems && p_ref && best_cost && best_idx && (rng || controls));
  NEED(x == nullptr || (x->c == c && x->connected && p->n_problems == 1 && x->mode == CAT_P2P_MODE_LOCKSTEP));
  int rc = bind(c);
  if (rc) return rc;
  CatNvtxRange range(c, x ? "cat_mpc_step_p2p (host)" : "cat_mpc_step (host)");
  const size_t np = (size_t)p->n_problems;
  const size_t esz = mpc_control_size<T>(p->flags);
  const size_t n_ctl = rng ? 0 : np * (size_t)p->H * 3 * (size_t)p->K;
  CatMpcProblem* d_prob;
  unsigned char* d_ctl = nullptr;
  double* d_ref;
  double* d_bc;
  long long* d_bi;
  T* d_costs = nullptr;
  // Few problems (the sampling-MPC step proper): problems and reference rows go up in ONE copy from pinned staging,
  // and the kernel's last block writes the winners into device-mapped pinned memory, so the call makes no D2H
  // copy and no pageable transfer (each of which costs a driver round trip of ~10 us on a 0.1 ms step).
  const size_t b_prob = (np * sizeof(CatMpcProblem) + 15) & ~(size_t)15;
  const size_t b_ref = np * (size_t)p->H * 3 * sizeof(double);
  const bool staged = np * 16 <= kStageWin && b_prob + b_ref <= kStageBytes - kStageWin;
  if (!rng) SCRATCH(SLOT_B, n_ctl * esz, d_ctl);
  if (costs) SCRATCH(SLOT_E, np * (size_t)p->K * sizeof(T), d_costs);
  if (staged) {
    rc = stage(c);
    if (rc) return rc;
    unsigned char* d_up;
    SCRATCH(SLOT_A, b_prob + b_ref, d_up);
    std::memcpy(c->h_stage + kStageWin, problems, np * sizeof(CatMpcProblem));
    std::memcpy(c->h_stage + kStageWin + b_prob, p_ref, b_ref);
    CU(cudaMemcpyAsync(d_up, c->h_stage + kStageWin, b_prob + b_ref, cudaMemcpyHostToDevice, c->stream));
    d_prob = (CatMpcProblem*)d_up;
    d_ref = (double*)(d_up + b_prob);
    d_bc = (double*)c->d_stage;
  } else {
    SCRATCH(SLOT_A, np * sizeof(CatMpcProblem), d_prob);
    SCRATCH(SLOT_C, b_ref, d_ref);
    SCRATCH(SLOT_D, np * 16, d_bc);
    CU(cudaMemcpyAsync(d_prob, problems, np * sizeof(CatMpcProblem), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_ref, p_ref, b_ref, cudaMemcpyHostToDevice, c->stream));
  }
  d_bi = (long long*)(d_bc + np);
  CatMpcDev<T> q;
  rc = mpc_fill<T>(c, p, q);
  if (rc) return rc;
  if (hp) {
    rc = mpc_attach_programs<T>(c, hp, q);
    if (rc) return rc;
  }
  // the shared scratch slots serve the host call (it synchronises before it returns); a workspace selected for
  // *_dev launches on other streams is left alone
  // Streamed upload (one problem, explicit SoA candidates, shipped equations, a grid that is resident at once): ONE
  // rollout is launched right away and the candidates follow over PCIe in row groups of whole horizon steps, each
  // followed by a stream write of this call's epoch into the group's arrival word; the kernel reads the controls of a
  // step only after it has seen the word (cat_arrive_need).  The rollout then trails the copy by one step, where the
  // chunked form below ends with the whole rollout of its last chunk (~35 us: 50 serial steps at one warp per
  // scheduler).  Config 3 on one B200: 0.794 (two chunks; 0.816 with the eight of round 1) -> 0.773 ms (float32 on
  // the wire), 1.505 -> 1.482 ms (fp64); at most 1 GB of candidates, so that a group cannot outlast the kernel's 2 s patience.
  const bool aos_ctl = (p->flags & CAT_MPC_CONTROLS_AOS) != 0;
  const bool shipped_eq = q.eq_theta == CAT_EQ_DTHETA_SHIPPED && q.eq_gamma == CAT_EQ_DGAMMA_SHIPPED;
  // Rows must start on 128-byte lines: a line that straddles two row groups could be fetched (into L1 by the fp32
  // kernel's cp.async) before its second half has arrived.
  bool streamed = c->streamed && !rng && np == 1 && p->K >= 32768 && !aos_ctl && !hp && shipped_eq && staged &&
                  ((size_t)p->K * esz) % 128 == 0 && n_ctl * esz <= ((size_t)1 << 30) &&
                  nblk(p->K, mpc_block_size(c, p->K, 1)) <= (unsigned)c->sm_count * CAT_MPC_MINB;
  if (streamed) {
    // page-locked candidates only: a copy from pageable memory blocks the host while it is staged, so the rollout --
    // launched after the copies are enqueued -- would start when everything is already on the device; the chunked
    // form launches between its copies and keeps the overlap
    cudaPointerAttributes pa;
    if (cudaPointerGetAttributes(&pa, controls) != cudaSuccess) {
      cudaGetLastError();
      streamed = false;
    } else if (pa.type != cudaMemoryTypeHost) {
      streamed = false;
    }
  }
  unsigned int* d_arrive = nullptr;
  if (streamed) {
    void* pp = nullptr;
    rc = scratch(c, SLOT_ARRIVE, 64 * sizeof(unsigned int) + sizeof(CatArrivePlan), &pp, true);
    if (rc) return rc;
    d_arrive = (unsigned int*)pp;
  }
  unsigned int* const h_epoch = (unsigned int*)(c->h_stage ? c->h_stage + 4000 : nullptr);
  unsigned int* const h_err = (unsigned int*)(c->h_stage ? c->h_stage + 4008 : nullptr);
  void* const ws_saved = c->ws;
  c->ws = nullptr;
  if (streamed) {
    CatNvtxRange r_up(c, "cat: streamed H2D of the candidates under the rollout");
    // Row groups.  Every boundary between two copies costs ~9 us on the link (measured: profiles/r2_ab_streamed_upload.txt),
    // a step of the rollout ~1.7 us and its row 15-30 us of copy, so the groups shrink towards the end by about the
    // ratio of the two: [the rest, H/10 steps, 1 step] -- the rollout of each group finishes under the next copy and
    // one step is left when the last byte lands.
    const int H = p->H;
    int sizes[CAT_ARRIVE_MAX] = {0}, ng = 0, sum = 0;
    for (int i = 0; i < CAT_ARRIVE_MAX && c->stream_split[i] > 0; ++i) sum += (sizes[ng++] = c->stream_split[i]);
    if (ng == 0 || sum != H) {
      const int last = H >= 2 ? 1 : 0, mid = H >= 12 ? (H + 5) / 10 : 0;
      ng = 0;
      sizes[ng++] = H - mid - last;
      if (mid) sizes[ng++] = mid;
      if (last) sizes[ng++] = last;
    }
    if (++c->arrive_epoch == 0) ++c->arrive_epoch;         // 0 is what freshly allocated words hold
    *h_epoch = c->arrive_epoch;
    *h_err = 0;
    // the plan rides behind the problem record and the reference rows in the staged upload (one more copy of 56 bytes,
    // ordered ahead of the rollout on its stream)
    CatArrivePlan plan;
    plan.epoch = c->arrive_epoch;
    plan.n_groups = ng;
    for (int i = 0, end = 0; i < CAT_ARRIVE_MAX; ++i) plan.end[i] = i < ng ? (end += sizes[i]) : H;
    plan.words = d_arrive;
    plan.err = (unsigned int*)(c->d_stage + 4008);
    std::memcpy(c->h_stage + 3840, &plan, sizeof(plan));
    CatArrivePlan* d_plan = reinterpret_cast<CatArrivePlan*>(d_arrive + 64);
    cudaError_t e = cudaMemcpyAsync(d_plan, c->h_stage + 3840, sizeof(plan), cudaMemcpyHostToDevice, c->stream);
    q.arrive = d_plan;
    const size_t row = (size_t)p->K * esz * 3;             // bytes of one horizon step of [H][3][K]
    const unsigned char* h_ctl = (const unsigned char*)controls;
    for (int g = 0, h0 = 0; g < ng && e == cudaSuccess; ++g) {
      const int h1 = h0 + sizes[g];
      e = cudaMemcpyAsync(d_ctl + (size_t)h0 * row, h_ctl + (size_t)h0 * row, (size_t)(h1 - h0) * row,
                          cudaMemcpyHostToDevice, c->aux_stream);
      if (e == cudaSuccess) {
        typedef int (*WriteValue32)(cudaStream_t, unsigned long long, unsigned int, unsigned int);
        // a device or driver without stream memory operations refuses the call: 4-byte copies from then on
        if (c->write_value32 &&
            ((WriteValue32)c->write_value32)(c->aux_stream, (unsigned long long)(uintptr_t)(d_arrive + g),
                                             c->arrive_epoch, 0u) != 0)
          c->write_value32 = nullptr;
        if (!c->write_value32)
          e = cudaMemcpyAsync(d_arrive + g, h_epoch, sizeof(unsigned int), cudaMemcpyHostToDevice, c->aux_stream);
      }
      h0 = h1;
    }
    if (e != cudaSuccess) rc = fail(c, CAT_ECUDA, "streamed control upload", e);
    if (rc == CAT_OK) {
      if (x) p2p_arm(x, q.p2p);
      rc = mpc_launch<T>(c, q, d_prob, d_ctl, d_ref, d_bc, d_bi, d_costs, nullptr, nullptr);
      if (rc == CAT_OK && x) ++x->seq;
    }
  } else if (!rng && np == 1 && p->K >= 32768) {
    // One problem, explicit controls (config 3: 78.6 MB in fp64, 39.3 MB as float): the step is PCIe-bound, so
    // the candidates go up in chunks on the aux stream and chunk i's rollout overlaps chunk i+1's copy; the chunk
    // winners fold on the device in candidate order (same total order as one launch: lowest cost, then lowest
    // global index).
    // Chunk boundaries.  The step ends one whole rollout after the last copy, and a rollout of up to one block per SM
    // takes the same ~35 us (50 serial steps at one warp per scheduler) however few candidates it has; every extra
    // copy costs ~9 us of link time (profiles/r2_ab_streamed_upload.txt).  So: a last chunk of one block per SM, and
    // the rest in `chunk_n` equal chunks (1: its rollout runs under the last copy, which is 9x longer).
    const int NC = 8;
    CatNvtxRange r_up(c, "cat: chunked H2D of the candidates + rollouts");
    long long* d_pairs;
    SCRATCH(SLOT_F, (size_t)NC * 16, d_pairs);
    const bool aos = (p->flags & CAT_MPC_CONTROLS_AOS) != 0;
    const long long K = p->K, H3 = 3LL * p->H;
    const unsigned char* h_ctl = (const unsigned char*)controls;
    long long kb[NC + 1];
    int nc = 0;
    {
      long long last = c->chunk_last > 0 ? c->chunk_last : (long long)c->sm_count * 128;
      last = (last + 127) / 128 * 128;
      if (last > K / 2) last = (K / 2 + 127) / 128 * 128;
      const int n_head = c->chunk_n >= 1 && c->chunk_n < NC ? c->chunk_n : 1;
      const long long head = K - last;
      kb[0] = 0;
      for (int i = 1; i <= n_head; ++i) kb[i] = i == n_head ? head : (head * i / n_head + 127) / 128 * 128;
      kb[n_head + 1] = K;
      nc = n_head + 1;
    }
    int used = 0;
    for (int i = 0; i < nc && rc == CAT_OK; ++i) {
      const long long k0 = kb[i], k1 = kb[i + 1];
      if (k1 <= k0) continue;
      cudaError_t e;
      if (aos)
        e = cudaMemcpyAsync(d_ctl + (size_t)(k0 * H3) * esz, h_ctl + (size_t)(k0 * H3) * esz,
                            (size_t)(k1 - k0) * H3 * esz, cudaMemcpyHostToDevice, c->aux_stream);
      else
        e = cudaMemcpy2DAsync(d_ctl + (size_t)k0 * esz, (size_t)K * esz, h_ctl + (size_t)k0 * esz, (size_t)K * esz,
                              (size_t)(k1 - k0) * esz, (size_t)H3, cudaMemcpyHostToDevice, c->aux_stream);
      if (e == cudaSuccess) e = cudaEventRecord(c->ev_chunk[i], c->aux_stream);
      if (e == cudaSuccess) e = cudaStreamWaitEvent(c->stream, c->ev_chunk[i], 0);
      if (e != cudaSuccess) {
        rc = fail(c, CAT_ECUDA, "chunked control upload", e);
        break;
      }
      CatMpcDev<T> qi = q;
      qi.K = k1 - k0;
      qi.k_offset = p->k_offset + k0;
      rc = mpc_launch<T>(c, qi, d_prob, d_ctl + (size_t)(k0 * q.sk) * esz, d_ref, (double*)(d_pairs + 2 * used),
                         d_pairs + 2 * used + 1, d_costs ? d_costs + k0 : nullptr, nullptr, nullptr);
      ++used;
    }
    if (rc == CAT_OK) {
      if (x) {
        CatP2PDev d;
        p2p_arm(x, d);
        k_p2p_exchange<<<1, 64, 0, c->stream>>>(used, d_pairs, d, d_bc, d_bi);
      } else {
        k_reduce_winners<<<1, 32, 0, c->stream>>>(used, d_pairs, (long long*)d_bc);   // -> d_bc[0], d_bi[0]
      }
      ++c->launches;
      const cudaError_t e = cudaGetLastError();
      if (e != cudaSuccess) rc = fail(c, CAT_ECUDA, "kernel launch (winner fold)", e);
      else if (x) ++x->seq;
    }
  } else {
    if (!rng) {
      const cudaError_t e = cudaMemcpyAsync(d_ctl, controls, n_ctl * esz, cudaMemcpyHostToDevice, c->stream);
      if (e != cudaSuccess) rc = fail(c, CAT_ECUDA, "control upload", e);
    }
    if (rc == CAT_OK) {
      if (x) p2p_arm(x, q.p2p);
      rc = mpc_launch<T>(c, q, d_prob, d_ctl, d_ref, d_bc, d_bi, d_costs, nullptr, nullptr);
      if (rc == CAT_OK && x) ++x->seq;
    }
  }
  c->ws = ws_saved;
  if (rc) {
    cudaStreamSynchronize(c->aux_stream);
    cudaStreamSynchronize(c->stream);
    return rc;
  }
  if (!staged) {
    CU(cudaMemcpyAsync(best_cost, d_bc, np * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(best_idx, d_bi, np * sizeof(long long), cudaMemcpyDeviceToHost, c->stream));
  }
  if (costs)
    CU(cudaMemcpyAsync(costs, d_costs, np * (size_t)p->K * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
  {
    CatNvtxRange r_wait(c, x ? "cat: wait for the job-wide winner (in-kernel exchange)" : "cat: wait for the winner");
    CU(cudaStreamSynchronize(c->stream));
  }
  if (streamed && *h_err) {
    cudaStreamSynchronize(c->aux_stream);
    return fail(c, CAT_ECUDA, "streamed control upload: a row group did not reach the device within 2 s");
  }
  if (staged) {   // the winners are already in host memory (written through the mapping before the kernel ended)
    std::memcpy(best_cost, c->h_stage, np * sizeof(double));
    std::memcpy(best_idx, c->h_stage + np * sizeof(double), np * sizeof(long long));
  }
  return CAT_OK;
}

// ---------------------------------------------------------------------------------------------
// run-time equation programs (cat_kernels_expr.cuh)
int expr_validate(CatCtx* c, int32_t n_features, int32_t n_ops, const int32_t* ops, int32_t n_consts,
                  const double* consts) {
  NEED(n_features >= 1 && n_ops >= 1 && n_ops <= CAT_EXPR_MAX_OPS && ops && n_consts >= 0 && n_consts <= CAT_EXPR_MAX_OPS);
  NEED(n_consts == 0 || consts);
  int depth = 0;
  for (int i = 0; i < n_ops; ++i) {
    const int op = ops[2 * i], arg = ops[2 * i + 1];
    NEED(op >= 0 && op < CAT_OP_COUNT);
    if (op == CAT_OP_CONST) {
      NEED(arg >= 0 && arg < n_consts);
      ++depth;
    } else if (op == CAT_OP_VAR) {
      NEED(arg >= 0 && arg < n_features);
      ++depth;
    } else if (op == CAT_OP_ADD || op == CAT_OP_SUB || op == CAT_OP_MUL || op == CAT_OP_DIV || op == CAT_OP_POW) {
      NEED(depth >= 2);
      --depth;
    } else {
      NEED(depth >= 1);
      if (op == CAT_OP_POWI) NEED(arg >= -64 && arg <= 64);
    }
    NEED(depth <= CAT_EXPR_STACK);
  }
  NEED(depth == 1);
  return CAT_OK;
}

// program + optional scaler into SLOT_F: [ops int2 n_ops | consts double n_consts | mean T F | inv_scale T F]
template <typename T>
int expr_upload(CatCtx* c, int32_t F, int32_t n_ops, const int32_t* ops, int32_t n_consts, const double* consts,
                const double* mean, const double* scale, int2** d_ops, double** d_consts, T** d_mean, T** d_inv) {
  const size_t b_ops = (size_t)n_ops * sizeof(int2);
  const size_t b_c = (size_t)(n_consts ? n_consts : 1) * sizeof(double);
  const size_t b_s = (size_t)F * sizeof(T);
  unsigned char* base;
  SCRATCH(SLOT_F, b_ops + b_c + 2 * b_s + 64, base);
  *d_ops = (int2*)base;
  *d_consts = (double*)(base + b_ops);
  *d_mean = nullptr;
  *d_inv = nullptr;
  CU(cudaMemcpyAsync(*d_ops, ops, b_ops, cudaMemcpyHostToDevice, c->stream));
  if (n_consts) CU(cudaMemcpyAsync(*d_consts, consts, (size_t)n_consts * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  if (mean && scale) {
    std::vector<T> h(2 * (size_t)F);
    for (int j = 0; j < F; ++j) {
      h[j] = T(mean[j]);
      h[F + j] = T(1.0 / scale[j]);
    }
    *d_mean = (T*)(base + b_ops + b_c);
    *d_inv = *d_mean + F;
    CU(cudaMemcpyAsync(*d_mean, h.data(), 2 * b_s, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));   // h goes out of scope
  }
  return CAT_OK;
}

template <typename T>
int expr_eval_host(CatCtx* c, int64_t n, int32_t F, const T* X, int32_t n_ops, const int32_t* ops, int32_t n_consts,
                   const double* consts, const double* mean, const double* scale, T* out) {
  NEED(n >= 0 && (n == 0 || (X && out)) && ((mean == nullptr) == (scale == nullptr)));
  int rc = expr_validate(c, F, n_ops, ops, n_consts, consts);
  if (rc) return rc;
  if (n == 0) return CAT_OK;
  rc = bind(c);
  if (rc) return rc;
  T *dX, *dO, *dm, *di;
  int2* d_ops;
  double* d_c;
  SCRATCH(SLOT_A, (size_t)n * F * sizeof(T), dX);
  SCRATCH(SLOT_B, (size_t)n * sizeof(T), dO);
  rc = expr_upload<T>(c, F, n_ops, ops, n_consts, consts, mean, scale, &d_ops, &d_c, &dm, &di);
  if (rc) return rc;
  CU(cudaMemcpyAsync(dX, X, (size_t)n * F * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  const size_t smem = (size_t)n_ops * sizeof(int2) + (size_t)(n_consts + 1) * sizeof(T);
  k_expr_eval<T><<<nblk(n, 256), 256, smem, c->stream>>>(n, F, dX, d_ops, n_ops, d_c, n_consts, dm, di, dO);
  LAUNCHED();
  CU(cudaMemcpyAsync(out, dO, (size_t)n * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

template <typename T>
int expr_integrate_host(CatCtx* c, int64_t n_traj, int64_t T_len, int32_t F, const T* X, const T* time, const T* y0,
                        int32_t n_ops, const int32_t* ops, int32_t n_consts, const double* consts, int32_t method,
                        T* y_out, T* ydot_out) {
  NEED(n_traj >= 0 && T_len >= 0 && method >= 0 && method <= 2);
  int rc = expr_validate(c, F, n_ops, ops, n_consts, consts);
  if (rc) return rc;
  if (n_traj == 0 || T_len == 0) return CAT_OK;
  NEED(X && time && y0 && y_out);
  rc = bind(c);
  if (rc) return rc;
  T *dX, *dt, *dy0, *dinc, *dy, *dyd, *dm, *di;
  int2* d_ops;
  double* d_c;
  const size_t rows = (size_t)n_traj * T_len;
  SCRATCH(SLOT_A, rows * F * sizeof(T), dX);
  SCRATCH(SLOT_B, (size_t)T_len * sizeof(T), dt);
  SCRATCH(SLOT_C, (size_t)n_traj * sizeof(T), dy0);
  SCRATCH(SLOT_D, rows * sizeof(T), dinc);
  SCRATCH(SLOT_E, rows * sizeof(T), dy);
  SCRATCH(SLOT_G, rows * sizeof(T), dyd);
  rc = expr_upload<T>(c, F, n_ops, ops, n_consts, consts, nullptr, nullptr, &d_ops, &d_c, &dm, &di);
  if (rc) return rc;
  CU(cudaMemcpyAsync(dX, X, rows * F * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dt, time, (size_t)T_len * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dy0, y0, (size_t)n_traj * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  const size_t smem = (size_t)n_ops * sizeof(int2) + (size_t)(n_consts + 1) * sizeof(T);
  const long long work = (long long)n_traj * (T_len - 1);
  if (T_len > 1) {
    k_expr_increments<T><<<nblk(work, 256), 256, smem, c->stream>>>(n_traj, T_len, F, dX, dt, d_ops, n_ops, d_c,
                                                                     n_consts, method, dinc);
    LAUNCHED();
  }
  if (method == 2) {
    // test_cluster.py:110-129: ydot starts at 0, y at y0
    CU(cudaMemsetAsync(dy, 0, (size_t)n_traj * sizeof(T), c->stream));   // zero y0 of the first pass
    k_rk4_accumulate<T><<<nblk(n_traj, 128), 128, 0, c->stream>>>(n_traj, T_len, dy, dinc, dyd);
    LAUNCHED();
    if (T_len > 1) {
      k_expr_second_pass<T><<<nblk(work, 256), 256, 0, c->stream>>>(n_traj, T_len, dt, dyd, dinc);
      LAUNCHED();
    }
    if (ydot_out) CU(cudaMemcpyAsync(ydot_out, dyd, rows * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
  }
  k_rk4_accumulate<T><<<nblk(n_traj, 128), 128, 0, c->stream>>>(n_traj, T_len, dy0, dinc, dy);
  LAUNCHED();
  CU(cudaMemcpyAsync(y_out, dy, rows * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

// ---------------------------------------------------------------------------------------------
// receding-horizon controller: device-resident state, one CUDA graph per control step
struct CatCtrl {
  CatCtx* c;
  bool f32;
  CatMpcParams p;
  CatCtrlConfig cfg;
  int64_t ref_len;
  int steps;                 // control steps enqueued so far
  int slices;                // MPPI candidate slices per horizon step
  CatMpcProblem* d_prob;
  double *d_unom, *d_ustar, *d_ref, *d_log, *d_bc;
  long long* d_bi;
  void* d_costs;             // T [n_problems][K] (MPPI only)
  int* d_step;
  unsigned int *d_ticket, *d_tk;
  CatBest* d_bb;
  cudaGraphExec_t exec;
  cudaStream_t exec_stream;  // stream the graph was captured on
};

void ctrl_free(CatCtrl* k) {
  if (!k) return;
  cudaSetDevice(k->c->device);
  if (k->exec) cudaGraphExecDestroy(k->exec);
  void* bufs[] = {k->d_prob, k->d_unom, k->d_ustar, k->d_ref, k->d_log, k->d_bc, k->d_costs, k->d_step,
                  k->d_ticket, k->d_tk, k->d_bb};
  for (void* b : bufs)
    if (b) cudaFree(b);
  delete k;
}

template <typename T>
int ctrl_fill(CatCtrl* k, CatMpcDev<T>& q) {
  CatCtx* c = k->c;
  int rc = mpc_fill<T>(c, &k->p, q);
  if (rc) return rc;
  q.flags = CAT_MPC_RNG_CONTROLS;
  q.u_nom = k->d_unom;
  q.step = k->d_step;
  q.ref_len = (int)k->ref_len;
  for (int a = 0; a < 3; ++a) q.u_max[a] = T(k->cfg.u_max[a]);
  return CAT_OK;
}

// the kernels of ONE control step, enqueued on the ctx stream
template <typename T>
int ctrl_enqueue_step(CatCtrl* k) {
  CatCtx* c = k->c;
  CatMpcDev<T> q;
  CatMpcDev<double> qd;
  int rc = ctrl_fill<T>(k, q);
  if (rc) return rc;
  rc = ctrl_fill<double>(k, qd);
  if (rc) return rc;
  CatCtrlDev x;
  x.mode = k->cfg.mode;
  x.max_steps = k->cfg.max_steps;
  x.lambda = k->cfg.lambda;
  x.slices = k->slices;
  const bool mppi = k->cfg.mode == CAT_CTRL_MPPI;
  rc = mpc_launch<T>(c, q, k->d_prob, nullptr, k->d_ref, k->d_bc, k->d_bi, (T*)k->d_costs, k->d_bb, k->d_tk);
  if (rc) return rc;
  if (mppi) {
    k_ctrl_mppi<T><<<dim3((unsigned)q.H, (unsigned)q.n_problems, (unsigned)k->slices), 256, 0, c->stream>>>(
        q, x, (const T*)k->d_costs, k->d_bc, k->d_bi, k->d_ustar);
    LAUNCHED();
  }
  k_ctrl_advance<T><<<(unsigned)q.n_problems, 64, (size_t)q.H * 3 * sizeof(double), c->stream>>>(
      q, qd, x, k->d_prob, k->d_unom, k->d_ustar, k->d_ref, k->d_bc, k->d_bi, k->d_log, k->d_step, k->d_ticket);
  LAUNCHED();
  return CAT_OK;
}

int ctrl_create(CatCtx* c, bool f32, const CatMpcParams* p, const CatCtrlConfig* cfg, const CatMpcProblem* problems,
                const double* u_nom0, const double* p_ref_traj, int64_t ref_len, void** out) {
  NEED(out != nullptr);
  *out = nullptr;
  NEED(p && cfg && problems && p_ref_traj);
  NEED(p->K >= 1 && p->H >= 1 && p->H <= 4096 && p->n_problems >= 1 && p->n_problems <= 65535);
  NEED(p->k_offset == 0);
  NEED(ref_len >= 1 && ref_len < (1LL << 30));
  NEED(cfg->mode == CAT_CTRL_ARGMIN || cfg->mode == CAT_CTRL_MPPI);
  NEED(cfg->max_steps >= 1);
  NEED(cfg->mode != CAT_CTRL_MPPI || cfg->lambda > 0);
  int rc = bind(c);
  if (rc) return rc;
  {
    CatMpcDev<double> chk;
    rc = mpc_fill<double>(c, p, chk);
    if (rc) return rc;
  }
  CatCtrl* k = new (std::nothrow) CatCtrl();
  if (!k) return fail(c, CAT_ENOMEM, "controller allocation");
  std::memset((void*)k, 0, sizeof(*k));
  k->c = c;
  k->f32 = f32;
  k->p = *p;
  k->p.flags = CAT_MPC_RNG_CONTROLS;
  k->cfg = *cfg;
  k->ref_len = ref_len;
  const size_t np = (size_t)p->n_problems, H = (size_t)p->H;
  const size_t n_blocks = ((size_t)p->K + 31) / 32;   // upper bound for any block size >= 32
  // MPPI reduction grid: enough (h, problem, slice) blocks for ~4 per SM, at least 1024 candidates per slice
  {
    long long want = (4LL * c->sm_count + (long long)(np * H) - 1) / (long long)(np * H);
    const long long cap = (p->K + 1023) / 1024;
    if (want > cap) want = cap;
    if (want > 64) want = 64;
    k->slices = want < 1 ? 1 : (int)want;
  }
  const size_t esz = f32 ? sizeof(float) : sizeof(double);
  bool ok = true;
  auto dm = [&](void** ptr, size_t bytes) {
    if (ok && cudaMalloc(ptr, bytes ? bytes : 16) != cudaSuccess) {
      ok = false;
      *ptr = nullptr;
    }
  };
  dm((void**)&k->d_prob, np * sizeof(CatMpcProblem));
  dm((void**)&k->d_unom, np * H * 3 * sizeof(double));
  dm((void**)&k->d_ustar, np * H * (size_t)k->slices * 4 * sizeof(double));
  dm((void**)&k->d_ref, np * (size_t)ref_len * 3 * sizeof(double));
  dm((void**)&k->d_log, np * (size_t)cfg->max_steps * CAT_CTRL_LOG_W * sizeof(double));
  dm((void**)&k->d_bc, np * 16);
  if (cfg->mode == CAT_CTRL_MPPI) dm(&k->d_costs, np * (size_t)p->K * esz);
  dm((void**)&k->d_step, 16);
  dm((void**)&k->d_ticket, 16);
  dm((void**)&k->d_tk, np * sizeof(unsigned int));
  dm((void**)&k->d_bb, np * n_blocks * sizeof(CatBest));
  if (!ok) {
    cudaGetLastError();
    ctrl_free(k);
    return fail(c, CAT_ENOMEM, "cudaMalloc (controller)");
  }
  k->d_bi = (long long*)(k->d_bc + np);
  cudaError_t e = cudaMemcpyAsync(k->d_prob, problems, np * sizeof(CatMpcProblem), cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess)
    e = u_nom0 ? cudaMemcpyAsync(k->d_unom, u_nom0, np * H * 3 * sizeof(double), cudaMemcpyHostToDevice, c->stream)
               : cudaMemsetAsync(k->d_unom, 0, np * H * 3 * sizeof(double), c->stream);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(k->d_ref, p_ref_traj, np * (size_t)ref_len * 3 * sizeof(double), cudaMemcpyHostToDevice,
                        c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(k->d_ustar, 0, np * H * (size_t)k->slices * 4 * sizeof(double), c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(k->d_step, 0, 16, c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(k->d_ticket, 0, 16, c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(k->d_tk, 0, np * sizeof(unsigned int), c->stream);
  if (e == cudaSuccess)
    e = cudaMemsetAsync(k->d_log, 0, np * (size_t)cfg->max_steps * CAT_CTRL_LOG_W * sizeof(double), c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);   // the host buffers may go away after this call
  if (e != cudaSuccess) {
    ctrl_free(k);
    return fail(c, CAT_ECUDA, "controller upload", e);
  }
  *out = k;
  return CAT_OK;
}

int ctrl_run(CatCtrl* k, int32_t n_steps) {
  CatCtx* c = k->c;
  NEED(n_steps >= 0 && (int64_t)k->steps + n_steps <= (int64_t)k->cfg.max_steps);
  int rc = bind(c);
  if (rc) return rc;
  if (n_steps == 0) return CAT_OK;
  if (k->cfg.use_graph) {
    if (k->exec && k->exec_stream != c->stream) {   // captured on another stream: rebuild
      cudaGraphExecDestroy(k->exec);
      k->exec = nullptr;
    }
    if (!k->exec) {
      cudaGraph_t g = nullptr;
      CU(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
      const int64_t before = c->launches;
      rc = k->f32 ? ctrl_enqueue_step<float>(k) : ctrl_enqueue_step<double>(k);
      c->launches = before;   // capturing is not launching
      cudaError_t e = cudaStreamEndCapture(c->stream, &g);
      if (rc) {
        if (g) cudaGraphDestroy(g);
        return rc;
      }
      if (e != cudaSuccess) return fail(c, CAT_ECUDA, "cudaStreamEndCapture", e);
      e = cudaGraphInstantiate(&k->exec, g, 0);
      cudaGraphDestroy(g);
      if (e != cudaSuccess) {
        k->exec = nullptr;
        return fail(c, CAT_ECUDA, "cudaGraphInstantiate", e);
      }
      k->exec_stream = c->stream;
    }
    const int per_step = k->cfg.mode == CAT_CTRL_MPPI ? 3 : 2;
    for (int i = 0; i < n_steps; ++i) {
      CU(cudaGraphLaunch(k->exec, c->stream));
      c->launches += per_step;
    }
  } else {
    for (int i = 0; i < n_steps; ++i) {
      rc = k->f32 ? ctrl_enqueue_step<float>(k) : ctrl_enqueue_step<double>(k);
      if (rc) return rc;
    }
  }
  k->steps += n_steps;
  return CAT_OK;
}

// ---------------------------------------------------------------------------------------------
// fused winner exchange over NVLink peer memory: device-pointer step
void p2p_free(CatP2P* x) {
  if (!x) return;
  cudaSetDevice(x->c->device);
  cudaStreamSynchronize(x->c->stream);
  for (int i = 0; i < x->n_ranks; ++i)
    if (x->opened[i]) cudaIpcCloseMemHandle(x->opened[i]);
  if (x->slots) cudaFree(x->slots);
  if (x->d_peers) cudaFree(x->d_peers);
  if (x->d_err) cudaFree(x->d_err);
  delete x;
}

template <typename T>
int mpc_step_p2p_dev(CatCtx* c, CatP2P* x, const CatMpcParams* p, const CatMpcProblem* problems,
                     const void* controls, const double* p_ref, double* best_cost, long long* best_idx, T* costs) {
  NEED(x != nullptr && x->c == c && x->connected);
  CatMpcDev<T> q;
  int rc = mpc_fill<T>(c, p, q);
  if (rc) return rc;
  const bool rng = (p->flags & CAT_MPC_RNG_CONTROLS) != 0;
  NEED(p->n_problems == 1 && problems && p_ref && best_cost && best_idx && (rng || controls));
  CatNvtxRange r(c, x->mode == CAT_P2P_MODE_PIPELINED ? "cat_mpc_step_p2p_dev (pipelined)" : "cat_mpc_step_p2p_dev");
  p2p_arm(x, q.p2p);
  rc = mpc_launch<T>(c, q, problems, controls, p_ref, best_cost, best_idx, costs, nullptr, nullptr);
  if (rc == CAT_OK) ++x->seq;   // a launch that failed posts nothing: the ranks stay in step
  return rc;
}

}  // namespace

#define CTX()                          \
  CatCtx* c = (CatCtx*)ctx;            \
  if (!c) return CAT_EINVAL

// =============================================================================================
extern "C" {

int cat_abi_version(void) { return CAT_ABI_VERSION; }

int cat_init(int device, void** ctx) {
  if (!ctx) return CAT_EINVAL;
  *ctx = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return CAT_ENODEV;
  }
  if (device < 0 || device >= n) return CAT_EINVAL;
  CatCtx* c = new (std::nothrow) CatCtx();
  if (!c) return CAT_ENOMEM;
  c->device = device;
  c->launches = 0;
  c->mpc_block = 0;
  c->h_stage = c->d_stage = nullptr;
  c->ws = nullptr;
  c->overlap = 1;
  if (const char* e = std::getenv("CAT_MPC_OVERLAP")) c->overlap = std::atoi(e) != 0;
  c->n_reg = 0;
  c->nvtx = 0;
  if (const char* e = std::getenv("CAT_NVTX")) c->nvtx = std::atoi(e) != 0;
  c->streamed = 1;
  if (const char* e = std::getenv("CAT_MPC_STREAMED")) c->streamed = std::atoi(e) != 0;
  c->arrive_epoch = 0;
  c->chunk_n = 1;
  c->chunk_last = 0;
  if (const char* e = std::getenv("CAT_MPC_CHUNK_N")) c->chunk_n = std::atoi(e);
  if (const char* e = std::getenv("CAT_MPC_CHUNK_LAST")) c->chunk_last = std::atoll(e);
  for (int i = 0; i < 8; ++i) c->stream_split[i] = 0;
  if (const char* e = std::getenv("CAT_MPC_STREAM_SPLIT")) {
    int i = 0;
    for (const char* s = e; *s && i < 8; ++i) {
      c->stream_split[i] = std::atoi(s);
      while (*s && *s != ',') ++s;
      if (*s == ',') ++s;
    }
  }
  // arrival words are written with a stream memory operation (a semaphore release in the command stream) rather than
  // a 4-byte copy, which costs a DMA set-up between every two row groups; resolved at run time, nothing is linked
  c->write_value32 = nullptr;
  {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &fn, cudaEnableDefault, &qr) == cudaSuccess &&
        qr == cudaDriverEntryPointSuccess)
      c->write_value32 = fn;
    else
      cudaGetLastError();
    if (const char* e = std::getenv("CAT_MPC_STREAM_MEMOP"))
      if (std::atoi(e) == 0) c->write_value32 = nullptr;
  }
  if (const char* e = std::getenv("CAT_MPC_BLOCK_SIZE")) {
    const int v = std::atoi(e);
    if (v == 32 || v == 64 || v == 128) c->mpc_block = v;
  }
  std::memset(c->slot, 0, sizeof(c->slot));
  std::memset(c->cap, 0, sizeof(c->cap));
  bool ok = cudaSetDevice(device) == cudaSuccess;
  ok = ok && cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaStreamCreateWithFlags(&c->aux_stream, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaEventCreate(&c->ev0) == cudaSuccess && cudaEventCreate(&c->ev1) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&c->ev_aux, cudaEventDisableTiming) == cudaSuccess;
  for (int i = 0; i < 8; ++i) ok = ok && cudaEventCreateWithFlags(&c->ev_chunk[i], cudaEventDisableTiming) == cudaSuccess;
  ok = ok && cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device) == cudaSuccess;
  if (!ok) {
    cudaGetLastError();
    delete c;
    return CAT_ECUDA;
  }
  c->stream = c->own_stream;
  *ctx = c;
  return CAT_OK;
}

int cat_destroy(void* ctx) {
  CTX();
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->own_stream);
  cudaStreamSynchronize(c->aux_stream);
  for (int s = 0; s < N_SLOTS; ++s)
    if (c->slot[s]) cudaFree(c->slot[s]);
  if (c->h_stage) cudaFreeHost(c->h_stage);
  for (int i = 0; i < c->n_reg; ++i) cudaHostUnregister(c->reg[i].p);
  cudaEventDestroy(c->ev0);
  cudaEventDestroy(c->ev1);
  cudaEventDestroy(c->ev_aux);
  for (int i = 0; i < 8; ++i) cudaEventDestroy(c->ev_chunk[i]);
  cudaStreamDestroy(c->own_stream);
  cudaStreamDestroy(c->aux_stream);
  delete c;
  return CAT_OK;
}

int cat_sync(void* ctx) {
  CTX();
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

const char* cat_last_error(void* ctx) {
  CatCtx* c = (CatCtx*)ctx;
  return c ? c->err.c_str() : "null context";
}

int64_t cat_launch_count(void* ctx) {
  CatCtx* c = (CatCtx*)ctx;
  return c ? c->launches : -1;
}

void* cat_stream(void* ctx) {
  CatCtx* c = (CatCtx*)ctx;
  return c ? (void*)c->stream : nullptr;
}

int cat_set_stream(void* ctx, void* cuda_stream, int own) {
  CTX();
  c->stream = own ? c->own_stream : (cudaStream_t)cuda_stream;
  return CAT_OK;
}

int cat_host_alloc(void* ctx, size_t bytes, void** out) {
  CTX();
  NEED(out != nullptr);
  int rc = bind(c);
  if (rc) return rc;
  cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 16, cudaHostAllocMapped | cudaHostAllocPortable);   // device-visible under UVA
  if (e != cudaSuccess) return fail(c, CAT_ENOMEM, "cudaHostAlloc", e);
  return CAT_OK;
}

int cat_host_free(void* ctx, void* p) {
  CTX();
  if (p) CU(cudaFreeHost(p));
  return CAT_OK;
}

int cat_host_register(void* ctx, void* p, size_t bytes) {
  CTX();
  NEED(p != nullptr && bytes > 0);
  for (int i = 0; i < c->n_reg; ++i)
    if (c->reg[i].p == p) return c->reg[i].bytes >= bytes ? CAT_OK : fail(c, CAT_EINVAL, "range registered with a smaller size");
  NEED(c->n_reg < 64);
  int rc = bind(c);
  if (rc) return rc;
  cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(c, CAT_ECUDA, "cudaHostRegister", e);
  }
  c->reg[c->n_reg].p = p;
  c->reg[c->n_reg].bytes = bytes;
  ++c->n_reg;
  return CAT_OK;
}

int cat_host_unregister(void* ctx, void* p) {
  CTX();
  for (int i = 0; i < c->n_reg; ++i)
    if (c->reg[i].p == p) {
      cudaStreamSynchronize(c->stream);
      cudaStreamSynchronize(c->aux_stream);
      cudaError_t e = cudaHostUnregister(p);
      c->reg[i] = c->reg[--c->n_reg];
      if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(c, CAT_ECUDA, "cudaHostUnregister", e);
      }
      return CAT_OK;
    }
  return fail(c, CAT_EINVAL, "range was not registered on this context");
}

int cat_mpc_workspace_create(void* ctx, int64_t max_candidates, int32_t max_problems, void** ws) {
  CTX();
  NEED(ws != nullptr);
  *ws = nullptr;
  NEED(max_candidates >= 1 && max_problems >= 1 && max_problems <= 65535);
  int rc = bind(c);
  if (rc) return rc;
  CatMpcWs* w = new (std::nothrow) CatMpcWs();
  if (!w) return fail(c, CAT_ENOMEM, "workspace allocation");
  w->n_bb = (size_t)((max_candidates + 31) / 32) * (size_t)max_problems;   // any block size >= 32
  w->n_tk = (size_t)max_problems;
  w->bb = nullptr;
  w->tk = nullptr;
  cudaError_t e = cudaMalloc((void**)&w->bb, w->n_bb * sizeof(CatBest));
  if (e == cudaSuccess) e = cudaMalloc((void**)&w->tk, w->n_tk * sizeof(unsigned int));
  if (e == cudaSuccess) e = cudaMemset(w->tk, 0, w->n_tk * sizeof(unsigned int));   // synchronous: ready on any stream
  if (e != cudaSuccess) {
    cudaGetLastError();
    if (w->bb) cudaFree(w->bb);
    if (w->tk) cudaFree(w->tk);
    delete w;
    return fail(c, CAT_ENOMEM, "cudaMalloc (workspace)", e);
  }
  *ws = w;
  return CAT_OK;
}

int cat_mpc_workspace_destroy(void* ctx, void* ws) {
  CTX();
  CatMpcWs* w = (CatMpcWs*)ws;
  if (!w) return CAT_OK;
  if (c->ws == ws) c->ws = nullptr;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  cudaFree(w->bb);
  cudaFree(w->tk);
  delete w;
  return CAT_OK;
}

int cat_set_workspace(void* ctx, void* ws_or_null) {
  CTX();
  c->ws = ws_or_null;
  return CAT_OK;
}

int cat_set_overlap(void* ctx, int on) {
  CTX();
  c->overlap = on ? 1 : 0;
  return CAT_OK;
}

int cat_mpc_stats(void* ctx, int enable, int reset, uint64_t* out) {
  CTX();
  int rc = bind(c);
  if (rc) return rc;
  unsigned long long acc[4] = {0, 0, 0, 0};
  unsigned long long* o = out ? acc : nullptr;
  int e = cat_mpc_inst_stats<double, 0>(enable, reset, o);
  if (!e) e = cat_mpc_inst_stats<double, 1>(enable, reset, o);
  if (!e) e = cat_mpc_inst_stats<double, 2>(enable, reset, o);
  if (!e) e = cat_mpc_inst_stats<float, 0>(enable, reset, o);
  if (!e) e = cat_mpc_inst_stats<float, 1>(enable, reset, o);
  if (!e) e = cat_mpc_inst_stats<float, 2>(enable, reset, o);
  if (e) return fail(c, CAT_ECUDA, "cat_mpc_stats", (cudaError_t)e);
  if (out)
    for (int i = 0; i < 4; ++i) out[i] = acc[i];
  return CAT_OK;
}

int cat_timer_start(void* ctx) {
  CTX();
  CU(cudaEventRecord(c->ev0, c->stream));
  return CAT_OK;
}

int cat_timer_stop_ms(void* ctx, float* ms) {
  CTX();
  NEED(ms != nullptr);
  CU(cudaEventRecord(c->ev1, c->stream));
  CU(cudaEventSynchronize(c->ev1));
  CU(cudaEventElapsedTime(ms, c->ev0, c->ev1));
  return CAT_OK;
}

// ---- catenary parameter -----------------------------------------------------------------------
int cat_solve_param_f64(void* ctx, int64_t n, const double* l, const double* dH, const double* L,
                        int64_t L_stride, double* C_out) {
  CTX();
  return solve_param_host<double>(c, n, l, dH, L, L_stride, C_out);
}
int cat_solve_param_f32(void* ctx, int64_t n, const float* l, const float* dH, const float* L, int64_t L_stride,
                        float* C_out) {
  CTX();
  return solve_param_host<float>(c, n, l, dH, L, L_stride, C_out);
}
int cat_solve_param_dev_f64(void* ctx, int64_t n, const double* l, const double* dH, const double* L,
                            int64_t L_stride, double* C_out) {
  CTX();
  NEED(n >= 0 && (n == 0 || (l && dH && L && C_out)) && (L_stride == 0 || L_stride == 1));
  int rc = bind(c);
  if (rc) return rc;
  return solve_param_dev<double>(c, n, l, dH, L, L_stride, C_out);
}
int cat_solve_param_dev_f32(void* ctx, int64_t n, const float* l, const float* dH, const float* L,
                            int64_t L_stride, float* C_out) {
  CTX();
  NEED(n >= 0 && (n == 0 || (l && dH && L && C_out)) && (L_stride == 0 || L_stride == 1));
  int rc = bind(c);
  if (rc) return rc;
  return solve_param_dev<float>(c, n, l, dH, L, L_stride, C_out);
}

int cat_span_terms_f64(void* ctx, int64_t n, const double* l, const double* dH, const double* L, int64_t L_stride,
                       double w, double* C_out, double* x_low, double* z_low, double* tension) {
  CTX();
  NEED(n >= 0 && (n == 0 || (l && dH && L)) && (L_stride == 0 || L_stride == 1));
  if (n == 0) return CAT_OK;
  int rc = bind(c);
  if (rc) return rc;
  const size_t nb = (size_t)n * sizeof(double);
  double *in, *o;
  SCRATCH(SLOT_A, 3 * nb, in);
  SCRATCH(SLOT_B, 4 * nb, o);
  CU(cudaMemcpyAsync(in, l, nb, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(in + n, dH, nb, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(in + 2 * n, L, L_stride ? nb : sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_span_terms<double><<<nblk(n, 256), 256, 0, c->stream>>>(n, in, in + n, in + 2 * n, L_stride, kCLo, kCHi, w,
                                                           C_out ? o : nullptr, x_low ? o + n : nullptr,
                                                           z_low ? o + 2 * n : nullptr,
                                                           tension ? o + 3 * n : nullptr);
  LAUNCHED();
  if (C_out) CU(cudaMemcpyAsync(C_out, o, nb, cudaMemcpyDeviceToHost, c->stream));
  if (x_low) CU(cudaMemcpyAsync(x_low, o + n, nb, cudaMemcpyDeviceToHost, c->stream));
  if (z_low) CU(cudaMemcpyAsync(z_low, o + 2 * n, nb, cudaMemcpyDeviceToHost, c->stream));
  if (tension) CU(cudaMemcpyAsync(tension, o + 3 * n, nb, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

int cat_rodrigues_f64(void* ctx, int64_t n, const double* v, const double* axis, int64_t axis_stride,
                      const double* angle, int64_t angle_stride, const double* origin, double* out) {
  CTX();
  NEED(n >= 0 && (n == 0 || (v && axis && angle && out)));
  NEED((axis_stride == 0 || axis_stride == 1) && (angle_stride == 0 || angle_stride == 1));
  if (n == 0) return CAT_OK;
  int rc = bind(c);
  if (rc) return rc;
  const size_t nb = (size_t)n * sizeof(double);
  double *dv, *dax, *dan, *dor, *dout;
  SCRATCH(SLOT_A, 3 * nb, dv);
  SCRATCH(SLOT_B, axis_stride ? 3 * nb : 24, dax);
  SCRATCH(SLOT_C, angle_stride ? nb : 8, dan);
  SCRATCH(SLOT_D, 24, dor);
  SCRATCH(SLOT_E, 3 * nb, dout);
  CU(cudaMemcpyAsync(dv, v, 3 * nb, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dax, axis, axis_stride ? 3 * nb : 24, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dan, angle, angle_stride ? nb : 8, cudaMemcpyHostToDevice, c->stream));
  if (origin) CU(cudaMemcpyAsync(dor, origin, 24, cudaMemcpyHostToDevice, c->stream));
  k_rodrigues<double><<<nblk(n, 256), 256, 0, c->stream>>>(n, dv, dax, axis_stride, dan, angle_stride,
                                                          origin ? dor : nullptr, dout);
  LAUNCHED();
  CU(cudaMemcpyAsync(out, dout, 3 * nb, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

// ---- shapes -------------------------------------------------------------------------------------
int cat_shape_f64(void* ctx, int64_t n, const double* A, const double* B, const double* L, int64_t L_stride,
                  const double* theta, const double* gamma, int32_t n_pts, int32_t which, double* out,
                  double* C_out, uint8_t* valid) {
  CTX();
  return shape_host<double>(c, n, A, B, L, L_stride, theta, gamma, n_pts, which, out, C_out, valid);
}
int cat_shape_f32(void* ctx, int64_t n, const float* A, const float* B, const float* L, int64_t L_stride,
                  const float* theta, const float* gamma, int32_t n_pts, int32_t which, float* out, float* C_out,
                  uint8_t* valid) {
  CTX();
  return shape_host<float>(c, n, A, B, L, L_stride, theta, gamma, n_pts, which, out, C_out, valid);
}
int cat_shape_dev_f64(void* ctx, int64_t n, const double* A, const double* B, const double* L, int64_t L_stride,
                      const double* theta, const double* gamma, int32_t n_pts, int32_t which, double* out,
                      double* C_out, uint8_t* valid) {
  CTX();
  NEED(n >= 0 && n_pts >= 1 && which > 0 && which < 16 && (L_stride == 0 || L_stride == 1));
  NEED(n == 0 || (A && B && L && out));
  int rc = bind(c);
  if (rc) return rc;
  return shape_dev_on<double>(c, c->stream, n, A, B, L, L_stride, theta, gamma, n_pts, which, out, C_out, valid);
}
int cat_shape_dev_f32(void* ctx, int64_t n, const float* A, const float* B, const float* L, int64_t L_stride,
                      const float* theta, const float* gamma, int32_t n_pts, int32_t which, float* out,
                      float* C_out, uint8_t* valid) {
  CTX();
  NEED(n >= 0 && n_pts >= 1 && which > 0 && which < 16 && (L_stride == 0 || L_stride == 1));
  NEED(n == 0 || (A && B && L && out));
  int rc = bind(c);
  if (rc) return rc;
  return shape_dev_on<float>(c, c->stream, n, A, B, L, L_stride, theta, gamma, n_pts, which, out, C_out, valid);
}

int cat_lowest_point_f64(void* ctx, int64_t n, int32_t n_pts, const double* curves, int64_t* idx_out,
                         double* pt_out) {
  CTX();
  NEED(n >= 0 && n_pts >= 1 && (n == 0 || (curves && idx_out && pt_out)));
  if (n == 0) return CAT_OK;
  int rc = bind(c);
  if (rc) return rc;
  double *dc, *dp;
  long long* di;
  const size_t nb = (size_t)n * n_pts * 3 * sizeof(double);
  SCRATCH(SLOT_A, nb, dc);
  SCRATCH(SLOT_B, (size_t)n * sizeof(long long), di);
  SCRATCH(SLOT_C, (size_t)n * 3 * sizeof(double), dp);
  CU(cudaMemcpyAsync(dc, curves, nb, cudaMemcpyHostToDevice, c->stream));
  k_lowest_point<double><<<nblk(n * 32, 256), 256, 0, c->stream>>>(n, n_pts, dc, di, dp);
  LAUNCHED();
  CU(cudaMemcpyAsync(idx_out, di, (size_t)n * sizeof(long long), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(pt_out, dp, (size_t)n * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

// ---- equations / integrators ----------------------------------------------------------------------
int cat_equation_count(void) { return CAT_N_EQUATIONS; }
int cat_equation_shipped(int which) {
  return which == 0 ? CAT_EQ_DTHETA_SHIPPED : which == 1 ? CAT_EQ_DGAMMA_SHIPPED : -1;
}
int cat_equation_info(int eq_id, int* is_gamma, int* complexity, uint32_t* live_mask) {
  if (eq_id < 0 || eq_id >= CAT_N_EQUATIONS) return CAT_EINVAL;
  if (is_gamma) *is_gamma = CAT_EQ_TABLE[eq_id].is_gamma;
  if (complexity) *complexity = CAT_EQ_TABLE[eq_id].complexity;
  if (live_mask) *live_mask = CAT_EQ_TABLE[eq_id].live_mask;
  return CAT_OK;
}

int cat_predict_f64(void* ctx, int64_t n, const double* X, int eq_id, int apply_scaler, double* out) {
  CTX();
  return predict_host<double>(c, n, X, eq_id, apply_scaler, out);
}
int cat_predict_f32(void* ctx, int64_t n, const float* X, int eq_id, int apply_scaler, float* out) {
  CTX();
  return predict_host<float>(c, n, X, eq_id, apply_scaler, out);
}
int cat_rk4_open_f64(void* ctx, int64_t n_traj, int64_t T, const double* X, const double* time, const double* y0,
                     int eq_id, int method, double* y_out) {
  CTX();
  return rk4_open_host<double>(c, n_traj, T, X, time, y0, eq_id, method, y_out);
}
int cat_rk4_open_f32(void* ctx, int64_t n_traj, int64_t T, const float* X, const float* time, const float* y0,
                     int eq_id, int method, float* y_out) {
  CTX();
  return rk4_open_host<float>(c, n_traj, T, X, time, y0, eq_id, method, y_out);
}

int cat_features_f64(void* ctx, int64_t T, const double* P0, const double* P1, const double* V1, const double* time,
                     const double* theta, const double* gamma, int apply_scaler, double* X_out) {
  CTX();
  NEED(T >= 0 && (T == 0 || (P0 && P1 && V1 && time && theta && gamma && X_out)));
  if (T == 0) return CAT_OK;
  int rc = bind(c);
  if (rc) return rc;
  double* in;   // P0(3) P1(3) V1(3) time theta gamma = 12 per row
  double* dX;
  SCRATCH(SLOT_A, (size_t)T * 12 * sizeof(double), in);
  SCRATCH(SLOT_B, (size_t)T * 18 * sizeof(double), dX);
  double *dP0 = in, *dP1 = in + 3 * T, *dV1 = in + 6 * T, *dt = in + 9 * T, *dth = in + 10 * T, *dga = in + 11 * T;
  CU(cudaMemcpyAsync(dP0, P0, (size_t)T * 24, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dP1, P1, (size_t)T * 24, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dV1, V1, (size_t)T * 24, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dt, time, (size_t)T * 8, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dth, theta, (size_t)T * 8, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dga, gamma, (size_t)T * 8, cudaMemcpyHostToDevice, c->stream));
  k_features<double><<<nblk(T, 256), 256, 0, c->stream>>>(T, dP0, dP1, dV1, dt, dth, dga, apply_scaler, dX);
  LAUNCHED();
  CU(cudaMemcpyAsync(X_out, dX, (size_t)T * 18 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

// ---- sampling MPC -------------------------------------------------------------------------------------
int cat_mpc_step_f64(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems, const double* controls,
                     const double* p_ref, double* best_cost, int64_t* best_idx, double* costs) {
  CTX();
  return mpc_step_host<double>(c, p, problems, controls, p_ref, best_cost, best_idx, costs);
}
int cat_mpc_step_f32(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems, const float* controls,
                     const double* p_ref, double* best_cost, int64_t* best_idx, float* costs) {
  CTX();
  return mpc_step_host<float>(c, p, problems, controls, p_ref, best_cost, best_idx, costs);
}
int cat_mpc_step_dev_f64(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems, const double* controls,
                         const double* p_ref, double* best_cost, int64_t* best_idx, double* costs) {
  CTX();
  int rc = bind(c);
  if (rc) return rc;
  return mpc_step_dev<double>(c, p, problems, controls, p_ref, best_cost, (long long*)best_idx, costs);
}
int cat_mpc_step_dev_f32(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems, const float* controls,
                         const double* p_ref, double* best_cost, int64_t* best_idx, float* costs) {
  CTX();
  int rc = bind(c);
  if (rc) return rc;
  return mpc_step_dev<float>(c, p, problems, controls, p_ref, best_cost, (long long*)best_idx, costs);
}

int cat_sample_controls_dev_f64(void* ctx, const CatMpcParams* p, double* controls) {
  CTX();
  CatMpcDev<double> q;
  int rc = mpc_fill<double>(c, p, q);
  if (rc) return rc;
  NEED(controls != nullptr && p->H <= 65535);
  rc = bind(c);
  if (rc) return rc;
  const dim3 grid(nblk(q.K, 256), (unsigned)q.H, (unsigned)q.n_problems);
  k_sample_controls<double><<<grid, 256, 0, c->stream>>>(q, controls);
  LAUNCHED();
  return CAT_OK;
}

int cat_reduce_winners_f64(int n, const double* cost, const int64_t* idx, double* best_cost, int64_t* best_idx) {
  if (n <= 0 || !cost || !idx || !best_cost || !best_idx) return CAT_EINVAL;
  double bc = cost[0];
  long long bi = idx[0];
  for (int g = 1; g < n; ++g)
    if (cat_better_mpc(cost[g], idx[g], bc, bi)) {
      bc = cost[g];
      bi = idx[g];
    }
  *best_cost = bc;
  *best_idx = bi;
  return CAT_OK;
}

int cat_reduce_winners_dev_f64(void* ctx, int n, const int64_t* gathered, int64_t* out) {
  CTX();
  NEED(n >= 1 && gathered && out);
  int rc = bind(c);
  if (rc) return rc;
  k_reduce_winners<<<1, 32, 0, c->stream>>>(n, (const long long*)gathered, (long long*)out);
  LAUNCHED();
  return CAT_OK;
}

// ---- roofline denominators --------------------------------------------------------------------------
int cat_catenary3d_f64(void* ctx, int64_t n, const double* p0, const double* p1, const double* L, int64_t L_stride,
                       int32_t n_pts, double* out) {
  CTX();
  NEED(n >= 0 && n_pts >= 1 && (L_stride == 0 || L_stride == 1));
  if (n == 0) return CAT_OK;
  NEED(p0 && p1 && L && out);
  int rc = bind(c);
  if (rc) return rc;
  double *d0, *d1, *dL, *da, *dout;
  const size_t b3 = (size_t)n * 3 * sizeof(double);
  SCRATCH(SLOT_A, b3, d0);
  SCRATCH(SLOT_B, b3, d1);
  SCRATCH(SLOT_C, (L_stride ? (size_t)n : 1) * sizeof(double), dL);
  SCRATCH(SLOT_D, (size_t)n * sizeof(double), da);
  SCRATCH(SLOT_E, (size_t)n * n_pts * 3 * sizeof(double), dout);
  CU(cudaMemcpyAsync(d0, p0, b3, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d1, p1, b3, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dL, L, (L_stride ? (size_t)n : 1) * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_catenary3d_solve<double><<<nblk(n, 256), 256, 0, c->stream>>>(n, d0, d1, dL, L_stride, da);
  LAUNCHED();
  k_catenary3d_points<double><<<nblk(n * (long long)n_pts, 256), 256, 0, c->stream>>>(n, n_pts, d0, d1, da, dout);
  LAUNCHED();
  CU(cudaMemcpyAsync(out, dout, (size_t)n * n_pts * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}
int cat_catenary_equation_f64(void* ctx, int64_t n, const double* x, const double* C, int64_t C_stride, double* y) {
  CTX();
  NEED(n >= 0 && (C_stride == 0 || C_stride == 1));
  if (n == 0) return CAT_OK;
  NEED(x && C && y);
  int rc = bind(c);
  if (rc) return rc;
  double *dx, *dC, *dy;
  SCRATCH(SLOT_A, (size_t)n * sizeof(double), dx);
  SCRATCH(SLOT_B, (C_stride ? (size_t)n : 1) * sizeof(double), dC);
  SCRATCH(SLOT_C, (size_t)n * sizeof(double), dy);
  CU(cudaMemcpyAsync(dx, x, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dC, C, (C_stride ? (size_t)n : 1) * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_catenary_equation<double><<<nblk(n, 256), 256, 0, c->stream>>>(n, dx, dC, C_stride, dy);
  LAUNCHED();
  CU(cudaMemcpyAsync(y, dy, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

int cat_resample_linear_f64(void* ctx, int64_t n_in, const double* time_in, int32_t n_cols, const double* cols,
                            int64_t n_out, const double* time_out, double* out) {
  CTX();
  NEED(n_in >= 2 && n_cols >= 1 && n_out >= 0 && time_in && cols);
  if (n_out == 0) return CAT_OK;
  NEED(time_out && out);
  int rc = bind(c);
  if (rc) return rc;
  double *dti, *dc, *dto, *dout;
  SCRATCH(SLOT_A, (size_t)n_in * sizeof(double), dti);
  SCRATCH(SLOT_B, (size_t)n_in * n_cols * sizeof(double), dc);
  SCRATCH(SLOT_C, (size_t)n_out * sizeof(double), dto);
  SCRATCH(SLOT_D, (size_t)n_out * n_cols * sizeof(double), dout);
  CU(cudaMemcpyAsync(dti, time_in, (size_t)n_in * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dc, cols, (size_t)n_in * n_cols * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dto, time_out, (size_t)n_out * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_resample_linear<double><<<nblk(n_out, 256), 256, 0, c->stream>>>(n_in, dti, n_cols, dc, n_out, dto, dout);
  LAUNCHED();
  CU(cudaMemcpyAsync(out, dout, (size_t)n_out * n_cols * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

int cat_feature_bank_f64(void* ctx, int64_t T, const double* P0, const double* P1, const double* V1,
                         const double* time, const double* theta, const double* gamma, double L, double w,
                         double* bank) {
  CTX();
  NEED(T >= 0 && L > 0);
  if (T == 0) return CAT_OK;
  NEED(P0 && P1 && V1 && time && theta && gamma && bank);
  int rc = bind(c);
  if (rc) return rc;
  double *d3, *d1, *dB;
  const size_t b3 = (size_t)T * 3 * sizeof(double), b1 = (size_t)T * sizeof(double);
  SCRATCH(SLOT_A, 3 * b3, d3);
  SCRATCH(SLOT_B, 3 * b1, d1);
  SCRATCH(SLOT_C, (size_t)T * CAT_BANK_W * sizeof(double), dB);
  CU(cudaMemcpyAsync(d3, P0, b3, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d3 + 3 * T, P1, b3, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d3 + 6 * T, V1, b3, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d1, time, b1, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d1 + T, theta, b1, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d1 + 2 * T, gamma, b1, cudaMemcpyHostToDevice, c->stream));
  k_feature_bank<double><<<nblk(T, 128), 128, 0, c->stream>>>(T, d3, d3 + 3 * T, d3 + 6 * T, d1, d1 + T, d1 + 2 * T, L,
                                                               w, kCLo, kCHi, dB);
  LAUNCHED();
  CU(cudaMemcpyAsync(bank, dB, (size_t)T * CAT_BANK_W * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

// ---- smoothed / second-order feature layouts (cat_kernels_feat.cuh) ----------------------------------------
int cat_fir_reflect_f64(void* ctx, int64_t n_sig, int64_t len, const double* x, int32_t n_taps, const double* taps,
                        double* out) {
  CTX();
  NEED(n_sig >= 0 && len >= 0 && n_taps >= 1 && (n_taps & 1) && n_taps <= 4097 && taps);
  if (n_sig == 0 || len == 0) return CAT_OK;
  NEED(x && out);
  int rc = bind(c);
  if (rc) return rc;
  const size_t nb = (size_t)n_sig * len * sizeof(double);
  double *dx, *dt, *dout;
  SCRATCH(SLOT_A, nb, dx);
  SCRATCH(SLOT_B, (size_t)n_taps * sizeof(double), dt);
  SCRATCH(SLOT_C, nb, dout);
  CU(cudaMemcpyAsync(dx, x, nb, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dt, taps, (size_t)n_taps * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_fir_reflect<double><<<nblk(n_sig * len, 256), 256, 0, c->stream>>>(n_sig, len, dx, n_taps, dt, dout);
  LAUNCHED();
  CU(cudaMemcpyAsync(out, dout, nb, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

int cat_savgol_f64(void* ctx, int64_t n_sig, int64_t len, const double* x, int32_t window, const double* taps,
                   const double* edge_l, const double* edge_r, double* out) {
  CTX();
  NEED(n_sig >= 0 && window >= 3 && (window & 1) && window <= 4097 && len >= window && taps && edge_l && edge_r);
  if (n_sig == 0) return CAT_OK;
  NEED(x && out);
  int rc = bind(c);
  if (rc) return rc;
  const size_t nb = (size_t)n_sig * len * sizeof(double);
  const size_t nw = (size_t)window, ne = (size_t)(window / 2) * window;
  double *dx, *dw, *dout;
  SCRATCH(SLOT_A, nb, dx);
  SCRATCH(SLOT_B, (nw + 2 * ne) * sizeof(double), dw);
  SCRATCH(SLOT_C, nb, dout);
  CU(cudaMemcpyAsync(dx, x, nb, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dw, taps, nw * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dw + nw, edge_l, ne * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dw + nw + ne, edge_r, ne * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_savgol<double><<<nblk(n_sig * len, 256), 256, 0, c->stream>>>(n_sig, len, dx, window, dw, dw + nw, dw + nw + ne, dout);
  LAUNCHED();
  CU(cudaMemcpyAsync(out, dout, nb, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

int cat_gradient_f64(void* ctx, int64_t n_sig, int64_t len, const double* y, const double* time, int32_t uniform,
                     double* out) {
  CTX();
  NEED(n_sig >= 0 && len >= 2 && time);
  if (n_sig == 0) return CAT_OK;
  NEED(y && out);
  int rc = bind(c);
  if (rc) return rc;
  const size_t nb = (size_t)n_sig * len * sizeof(double);
  double *dy, *dt, *dout;
  SCRATCH(SLOT_A, nb, dy);
  SCRATCH(SLOT_B, (size_t)len * sizeof(double), dt);
  SCRATCH(SLOT_C, nb, dout);
  CU(cudaMemcpyAsync(dy, y, nb, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dt, time, (size_t)len * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_gradient<double><<<nblk(n_sig * len, 256), 256, 0, c->stream>>>(n_sig, len, dy, dt, uniform, dout);
  LAUNCHED();
  CU(cudaMemcpyAsync(out, dout, nb, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

namespace {
// upload of an experiment's columns into one arena: [P0 (3T) | P1 (3T) | V1 (3T) | time (T) | theta (T) | gamma (T)]
int feat_upload(CatCtx* c, int64_t T, const double* P0, const double* P1, const double* V1, const double* time,
                const double* theta, const double* gamma, double** arena) {
  double* in;
  SCRATCH(SLOT_A, (size_t)T * 12 * sizeof(double), in);
  CU(cudaMemcpyAsync(in, P0, (size_t)T * 24, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(in + 3 * T, P1, (size_t)T * 24, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(in + 6 * T, V1, (size_t)T * 24, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(in + 9 * T, time, (size_t)T * 8, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(in + 10 * T, theta, (size_t)T * 8, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(in + 11 * T, gamma, (size_t)T * 8, cudaMemcpyHostToDevice, c->stream));
  *arena = in;
  return CAT_OK;
}
}  // namespace

int cat_features_valid_f64(void* ctx, int64_t T, const double* P0, const double* P1, const double* V1,
                           const double* time, const double* theta, const double* gamma, int32_t n_taps,
                           const double* taps, int32_t uniform, int32_t for_gamma, double* out) {
  CTX();
  NEED(T >= 0 && n_taps >= 1 && (n_taps & 1) && n_taps <= 4097 && taps);
  if (T == 0) return CAT_OK;
  NEED(T >= 2 && P0 && P1 && V1 && time && theta && gamma && out);
  int rc = bind(c);
  if (rc) return rc;
  CatNvtxRange range(c, "cat_features_valid");
  double *in, *w, *dt, *dout;
  rc = feat_upload(c, T, P0, P1, V1, time, theta, gamma, &in);
  if (rc) return rc;
  // work signals: sig (6T) | th1, ga1 (2T: the (T,1)-filtered angles) | sm (2T: smoothed) | d (2T: their gradient)
  SCRATCH(SLOT_B, (size_t)T * 12 * sizeof(double), w);
  SCRATCH(SLOT_C, (size_t)n_taps * sizeof(double), dt);
  SCRATCH(SLOT_D, (size_t)T * 13 * sizeof(double), dout);
  CU(cudaMemcpyAsync(dt, taps, (size_t)n_taps * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  double *sig = w, *a1 = w + 6 * T, *sm = w + 8 * T, *d = w + 10 * T;
  const double* ang = in + 10 * T;   // [theta | gamma]
  k_cable_terms<double><<<nblk(T, 256), 256, 0, c->stream>>>(T, in, in + 3 * T, in + 6 * T, 1.0, sig);
  LAUNCHED();
  // gaussian_filter1d(x.reshape(-1, 1)): the filtered axis has length 1 (main_fun.py:510-511)
  k_fir_reflect<double><<<nblk(2 * T, 256), 256, 0, c->stream>>>(2 * T, 1, ang, n_taps, dt, a1);
  LAUNCHED();
  k_fir_reflect<double><<<nblk(2 * T, 256), 256, 0, c->stream>>>(2, T, ang, n_taps, dt, sm);
  LAUNCHED();
  k_gradient<double><<<nblk(2 * T, 256), 256, 0, c->stream>>>(2, T, sm, in + 9 * T, uniform, d);
  LAUNCHED();
  k_valid_rows<double><<<nblk(T, 256), 256, 0, c->stream>>>(T, sig, a1, a1 + T, d, d + T, for_gamma, dout);
  LAUNCHED();
  CU(cudaMemcpyAsync(out, dout, (size_t)T * 13 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

int cat_features_dd_f64(void* ctx, int64_t T, const double* P0, const double* P1, const double* V1,
                        const double* time, const double* theta, const double* gamma, int32_t window,
                        const double* taps, const double* edge_l, const double* edge_r, int32_t uniform,
                        double* feats, double* targets) {
  CTX();
  NEED(T >= 0 && window >= 3 && (window & 1) && window <= 4097 && taps && edge_l && edge_r);
  if (T == 0) return CAT_OK;
  NEED(T >= window && P0 && P1 && V1 && time && theta && gamma && feats && targets);
  int rc = bind(c);
  if (rc) return rc;
  CatNvtxRange range(c, "cat_features_dd");
  double *in, *w, *dw, *dout;
  rc = feat_upload(c, T, P0, P1, V1, time, theta, gamma, &in);
  if (rc) return rc;
  // work signals: sig (6T) | sm (2T) | dsm (2T) | d2 (2T) | dsig (5T)
  SCRATCH(SLOT_B, (size_t)T * 17 * sizeof(double), w);
  const size_t nw = (size_t)window, ne = (size_t)(window / 2) * window;
  SCRATCH(SLOT_C, (nw + 2 * ne) * sizeof(double), dw);
  SCRATCH(SLOT_D, (size_t)T * 16 * sizeof(double), dout);
  CU(cudaMemcpyAsync(dw, taps, nw * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dw + nw, edge_l, ne * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(dw + nw + ne, edge_r, ne * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  double *sig = w, *sm = w + 6 * T, *dsm = w + 8 * T, *d2 = w + 10 * T, *dsig = w + 12 * T;
  const double* tm = in + 9 * T;
  k_cable_terms<double><<<nblk(T, 256), 256, 0, c->stream>>>(T, in, in + 3 * T, in + 6 * T, 1e-3, sig);   // V1 in m/s
  LAUNCHED();
  k_savgol<double><<<nblk(2 * T, 256), 256, 0, c->stream>>>(2, T, in + 10 * T, window, dw, dw + nw, dw + nw + ne, sm);
  LAUNCHED();
  k_gradient<double><<<nblk(2 * T, 256), 256, 0, c->stream>>>(2, T, sm, tm, uniform, dsm);
  LAUNCHED();
  k_gradient<double><<<nblk(2 * T, 256), 256, 0, c->stream>>>(2, T, dsm, tm, uniform, d2);
  LAUNCHED();
  k_gradient<double><<<nblk(5 * T, 256), 256, 0, c->stream>>>(5, T, sig + T, tm, uniform, dsig);
  LAUNCHED();
  k_dd_rows<double><<<nblk(T, 256), 256, 0, c->stream>>>(T, sig, sm, dsm, dsig, d2, dout, dout + 14 * T);
  LAUNCHED();
  CU(cudaMemcpyAsync(feats, dout, (size_t)T * 14 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(targets, dout + 14 * T, (size_t)T * 2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}

int cat_expr_eval_f64(void* ctx, int64_t n, int32_t n_features, const double* X, int32_t n_ops, const int32_t* ops,
                      int32_t n_consts, const double* consts, const double* mean, const double* scale, double* out) {
  CTX();
  return expr_eval_host<double>(c, n, n_features, X, n_ops, ops, n_consts, consts, mean, scale, out);
}
int cat_expr_eval_f32(void* ctx, int64_t n, int32_t n_features, const float* X, int32_t n_ops, const int32_t* ops,
                      int32_t n_consts, const double* consts, const double* mean, const double* scale, float* out) {
  CTX();
  return expr_eval_host<float>(c, n, n_features, X, n_ops, ops, n_consts, consts, mean, scale, out);
}
int cat_expr_integrate_f64(void* ctx, int64_t n_traj, int64_t T, int32_t n_features, const double* X,
                           const double* time, const double* y0, int32_t n_ops, const int32_t* ops,
                           int32_t n_consts, const double* consts, int32_t method, double* y_out,
                           double* ydot_out) {
  CTX();
  return expr_integrate_host<double>(c, n_traj, T, n_features, X, time, y0, n_ops, ops, n_consts, consts, method,
                                     y_out, ydot_out);
}

int cat_mpc_step_expr_f64(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems, const double* controls,
                          const double* p_ref, int32_t n_ops_theta, const int32_t* ops_theta, int32_t n_consts_theta,
                          const double* consts_theta, int32_t n_ops_gamma, const int32_t* ops_gamma,
                          int32_t n_consts_gamma, const double* consts_gamma, double* best_cost, int64_t* best_idx,
                          double* costs) {
  CTX();
  const CatHostPrograms hp = {n_ops_theta, n_consts_theta, n_ops_gamma, n_consts_gamma,
                              ops_theta,   ops_gamma,      consts_theta, consts_gamma};
  return mpc_step_host<double>(c, p, problems, controls, p_ref, best_cost, best_idx, costs, &hp);
}
int cat_mpc_step_expr_f32(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems, const float* controls,
                          const double* p_ref, int32_t n_ops_theta, const int32_t* ops_theta, int32_t n_consts_theta,
                          const double* consts_theta, int32_t n_ops_gamma, const int32_t* ops_gamma,
                          int32_t n_consts_gamma, const double* consts_gamma, double* best_cost, int64_t* best_idx,
                          float* costs) {
  CTX();
  const CatHostPrograms hp = {n_ops_theta, n_consts_theta, n_ops_gamma, n_consts_gamma,
                              ops_theta,   ops_gamma,      consts_theta, consts_gamma};
  return mpc_step_host<float>(c, p, problems, controls, p_ref, best_cost, best_idx, costs, &hp);
}

int cat_p2p_create(void* ctx, int rank, int n_ranks, void** p2p, void* handle_out) {
  CTX();
  NEED(p2p != nullptr);
  *p2p = nullptr;
  NEED(handle_out && n_ranks >= 1 && n_ranks <= CAT_P2P_MAX_RANKS && rank >= 0 && rank < n_ranks);
  int rc = bind(c);
  if (rc) return rc;
  CatP2P* x = new (std::nothrow) CatP2P();
  if (!x) return fail(c, CAT_ENOMEM, "p2p allocation");
  std::memset((void*)x, 0, sizeof(*x));
  x->c = c;
  x->rank = rank;
  x->n_ranks = n_ranks;
  x->mode = CAT_P2P_MODE_LOCKSTEP;
  const size_t bytes = (size_t)CAT_P2P_RING * n_ranks * CAT_P2P_SLOT_W * sizeof(long long);
  cudaError_t e = cudaMalloc((void**)&x->slots, bytes);
  if (e == cudaSuccess) e = cudaMemset(x->slots, 0, bytes);
  if (e == cudaSuccess) e = cudaMalloc((void**)&x->d_peers, (size_t)n_ranks * sizeof(long long*));
  if (e == cudaSuccess) e = cudaMalloc((void**)&x->d_err, 16);
  if (e == cudaSuccess) e = cudaMemset(x->d_err, 0, 16);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, x->slots);
  if (e != cudaSuccess) {
    p2p_free(x);
    return fail(c, CAT_ECUDA, "p2p buffers / cudaIpcGetMemHandle", e);
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  std::memcpy(handle_out, &h, 64);
  *p2p = x;
  return CAT_OK;
}
int cat_p2p_connect(void* p2p, const void* all_handles) {
  CatP2P* x = (CatP2P*)p2p;
  if (!x) return CAT_EINVAL;
  CatCtx* c = x->c;
  NEED(all_handles != nullptr && !x->connected);
  int rc = bind(c);
  if (rc) return rc;
  long long* table[CAT_P2P_MAX_RANKS];
  for (int r = 0; r < x->n_ranks; ++r) {
    if (r == x->rank) {
      table[r] = x->slots;
      continue;
    }
    cudaIpcMemHandle_t h;
    std::memcpy(&h, (const char*)all_handles + (size_t)64 * r, 64);
    void* ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return fail(c, CAT_ECUDA, "cudaIpcOpenMemHandle (peer slot buffer)", e);
    x->opened[r] = ptr;
    table[r] = (long long*)ptr;
  }
  CU(cudaMemcpy(x->d_peers, table, (size_t)x->n_ranks * sizeof(long long*), cudaMemcpyHostToDevice));
  x->connected = true;
  return CAT_OK;
}
int cat_mpc_step_p2p_dev_f64(void* ctx, void* p2p, const CatMpcParams* p, const CatMpcProblem* problems,
                             const double* controls, const double* p_ref, double* best_cost, int64_t* best_idx,
                             double* costs) {
  CTX();
  int rc = bind(c);
  if (rc) return rc;
  return mpc_step_p2p_dev<double>(c, (CatP2P*)p2p, p, problems, controls, p_ref, best_cost, (long long*)best_idx, costs);
}
int cat_mpc_step_p2p_dev_f32(void* ctx, void* p2p, const CatMpcParams* p, const CatMpcProblem* problems,
                             const float* controls, const double* p_ref, double* best_cost, int64_t* best_idx,
                             float* costs) {
  CTX();
  int rc = bind(c);
  if (rc) return rc;
  return mpc_step_p2p_dev<float>(c, (CatP2P*)p2p, p, problems, controls, p_ref, best_cost, (long long*)best_idx, costs);
}
int cat_mpc_step_p2p_f64(void* ctx, void* p2p, const CatMpcParams* p, const CatMpcProblem* problems,
                         const double* controls, const double* p_ref, double* best_cost, int64_t* best_idx,
                         double* costs) {
  CTX();
  NEED(p2p != nullptr);
  return mpc_step_host<double>(c, p, problems, controls, p_ref, best_cost, best_idx, costs, nullptr, (CatP2P*)p2p);
}
int cat_mpc_step_p2p_f32(void* ctx, void* p2p, const CatMpcParams* p, const CatMpcProblem* problems,
                         const float* controls, const double* p_ref, double* best_cost, int64_t* best_idx,
                         float* costs) {
  CTX();
  NEED(p2p != nullptr);
  return mpc_step_host<float>(c, p, problems, controls, p_ref, best_cost, best_idx, costs, nullptr, (CatP2P*)p2p);
}
int cat_p2p_set_mode(void* p2p, int mode) {
  CatP2P* x = (CatP2P*)p2p;
  if (!x) return CAT_EINVAL;
  CatCtx* c = x->c;
  NEED(mode == CAT_P2P_MODE_LOCKSTEP || mode == CAT_P2P_MODE_PIPELINED);
  x->mode = mode;
  return CAT_OK;
}
int cat_p2p_flush_dev(void* ctx, void* p2p, double* best_cost, int64_t* best_idx) {
  CTX();
  CatP2P* x = (CatP2P*)p2p;
  NEED(x != nullptr && x->c == c && x->connected && best_cost && best_idx);
  int rc = bind(c);
  if (rc) return rc;
  CatP2PDev d;
  p2p_arm(x, d);
  d.seq = x->seq;   // the step already launched, not the next one
  k_p2p_fold<<<1, 64, 0, c->stream>>>(d, best_cost, (long long*)best_idx);
  LAUNCHED();
  return CAT_OK;
}
int cat_p2p_status(void* p2p, int* timed_out) {
  CatP2P* x = (CatP2P*)p2p;
  if (!x || !timed_out) return CAT_EINVAL;
  CatCtx* c = x->c;
  int rc = bind(c);
  if (rc) return rc;
  CU(cudaStreamSynchronize(c->stream));
  CU(cudaMemcpy(timed_out, x->d_err, sizeof(int), cudaMemcpyDeviceToHost));
  if (*timed_out) CU(cudaMemset(x->d_err, 0, sizeof(int)));
  return CAT_OK;
}
int cat_p2p_destroy(void* p2p) {
  CatP2P* x = (CatP2P*)p2p;
  if (!x) return CAT_EINVAL;
  p2p_free(x);
  return CAT_OK;
}

int cat_ctrl_create_f64(void* ctx, const CatMpcParams* p, const CatCtrlConfig* cfg, const CatMpcProblem* problems,
                        const double* u_nom0, const double* p_ref_traj, int64_t ref_len, void** ctrl) {
  CTX();
  return ctrl_create(c, false, p, cfg, problems, u_nom0, p_ref_traj, ref_len, ctrl);
}
int cat_ctrl_create_f32(void* ctx, const CatMpcParams* p, const CatCtrlConfig* cfg, const CatMpcProblem* problems,
                        const double* u_nom0, const double* p_ref_traj, int64_t ref_len, void** ctrl) {
  CTX();
  return ctrl_create(c, true, p, cfg, problems, u_nom0, p_ref_traj, ref_len, ctrl);
}
int cat_ctrl_run(void* ctrl, int32_t n_steps) {
  CatCtrl* k = (CatCtrl*)ctrl;
  if (!k) return CAT_EINVAL;
  return ctrl_run(k, n_steps);
}
int cat_ctrl_set_state(void* ctrl, const CatMpcProblem* problems) {
  CatCtrl* k = (CatCtrl*)ctrl;
  if (!k) return CAT_EINVAL;
  CatCtx* c = k->c;
  NEED(problems != nullptr);
  int rc = bind(c);
  if (rc) return rc;
  CU(cudaMemcpyAsync(k->d_prob, problems, (size_t)k->p.n_problems * sizeof(CatMpcProblem), cudaMemcpyHostToDevice,
                     c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return CAT_OK;
}
int cat_ctrl_read(void* ctrl, int32_t* step_out, CatMpcProblem* problems_out, double* u_nom_out, double* log_out,
                  int32_t log_from, int32_t log_n) {
  CatCtrl* k = (CatCtrl*)ctrl;
  if (!k) return CAT_EINVAL;
  CatCtx* c = k->c;
  NEED(!log_out || (log_from >= 0 && log_n >= 0 && (int64_t)log_from + log_n <= (int64_t)k->cfg.max_steps));
  int rc = bind(c);
  if (rc) return rc;
  const size_t np = (size_t)k->p.n_problems, H = (size_t)k->p.H;
  int step = 0;
  if (step_out) CU(cudaMemcpyAsync(&step, k->d_step, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  if (problems_out)
    CU(cudaMemcpyAsync(problems_out, k->d_prob, np * sizeof(CatMpcProblem), cudaMemcpyDeviceToHost, c->stream));
  if (u_nom_out)
    CU(cudaMemcpyAsync(u_nom_out, k->d_unom, np * H * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  if (log_out && log_n > 0)
    CU(cudaMemcpy2DAsync(log_out, (size_t)log_n * CAT_CTRL_LOG_W * sizeof(double),
                         k->d_log + (size_t)log_from * CAT_CTRL_LOG_W,
                         (size_t)k->cfg.max_steps * CAT_CTRL_LOG_W * sizeof(double),
                         (size_t)log_n * CAT_CTRL_LOG_W * sizeof(double), np, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  if (step_out) *step_out = step;
  return CAT_OK;
}
int cat_ctrl_destroy(void* ctrl) {
  CatCtrl* k = (CatCtrl*)ctrl;
  if (!k) return CAT_EINVAL;
  cudaSetDevice(k->c->device);
  cudaStreamSynchronize(k->c->stream);
  ctrl_free(k);
  return CAT_OK;
}

int cat_measure_peak(void* ctx, int kind, double* ops_per_s) {
  CTX();
  NEED(ops_per_s != nullptr && kind >= 0 && kind <= 2);
  int rc = bind(c);
  if (rc) return rc;
  double* sink;
  SCRATCH(SLOT_A, 64, sink);
  const int blocks = c->sm_count * 8, threads = 256, iters = 4096;
  float best_ms = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CU(cudaEventRecord(c->ev0, c->stream));
    if (kind == 0)
      k_peak<0><<<blocks, threads, 0, c->stream>>>(iters, sink);
    else if (kind == 1)
      k_peak<1><<<blocks, threads, 0, c->stream>>>(iters, sink);
    else
      k_peak<2><<<blocks, threads, 0, c->stream>>>(iters, sink);
    LAUNCHED();
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaEventSynchronize(c->ev1));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    if (rep > 0 && ms < best_ms) best_ms = ms;
  }
  const double ops = (double)blocks * threads * (double)iters * 64.0;
  *ops_per_s = ops / (best_ms * 1e-3);
  return CAT_OK;
}

}  // extern "C"

// =============================================================================================
// Allocation-free sampling-MPC step (cat_mpc_step_static_dev_f64).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
// harnesses that forbid it).  The rollout's scratch -- one winner slot per thread block and one ticket per problem --
// lives in a file-scope device array of fixed size instead of the context's grow-on-demand slots; every GPU that
// loads the library holds its 528 KB until exit.  Every call on a device shares the array: calls that overlap on
// the GPU (two streams, two contexts) race on it, so they must be ordered by the caller.
// =============================================================================================
#define CAT_STR_(x) #x
#define CAT_STR(x) CAT_STR_(x)
constexpr int kStaticMaxBlocks = CAT_STATIC_MAX_BLOCKS;       // winner slots: blocks of 128 candidates x problems
constexpr int kStaticMaxProblems = CAT_STATIC_MAX_PROBLEMS;   // tickets
constexpr size_t kStaticTicketsAt = (size_t)kStaticMaxBlocks * sizeof(CatBest);
constexpr size_t kStaticBytes = kStaticTicketsAt + (size_t)kStaticMaxProblems * sizeof(unsigned int);
static __device__ __align__(256) unsigned char cat_static_scratch[kStaticBytes];

namespace {
// The array's name is a host-side stand-in in host code, not a device address: the address is resolved once per
// device, with that device current, and cached.
std::mutex g_static_mu;
unsigned char* g_static_addr[64] = {};

int static_scratch_addr(CatCtx* c, unsigned char** out) {
  NEED(c->device >= 0 && c->device < 64);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_addr[c->device]) {
    void* p = nullptr;
    CU(cudaGetSymbolAddress(&p, cat_static_scratch));   // the caller has made c->device current (bind)
    if (!p) return fail(c, CAT_ECUDA, "cudaGetSymbolAddress(cat_static_scratch) returned null");
    g_static_addr[c->device] = (unsigned char*)p;
  }
  *out = g_static_addr[c->device];
  return CAT_OK;
}
}  // namespace

extern "C" int cat_mpc_step_static_dev_f64(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems,
                                           const double* controls, const double* p_ref, double* best_cost,
                                           int64_t* best_idx, double* costs) {
  CTX();
  int rc = bind(c);
  if (rc) return rc;
  CatMpcDev<double> q;
  rc = mpc_fill<double>(c, p, q);
  if (rc) return rc;
  const bool rng = (p->flags & CAT_MPC_RNG_CONTROLS) != 0;
  NEED(problems && p_ref && best_cost && best_idx && (rng || controls));
  const long long n_blocks = (long long)nblk(q.K, mpc_block_size(c, q.K, q.n_problems)) * q.n_problems;
  if (n_blocks > kStaticMaxBlocks || q.n_problems > kStaticMaxProblems)
    return fail(c, CAT_EINVAL,
                "cat_mpc_step_static_dev_f64: step exceeds the static scratch (at most " CAT_STR(CAT_STATIC_MAX_BLOCKS)
                " thread blocks of 128 candidates summed over the problems, at most " CAT_STR(CAT_STATIC_MAX_PROBLEMS)
                " problems); use cat_mpc_step_dev_f64");
  unsigned char* base = nullptr;
  rc = static_scratch_addr(c, &base);
  if (rc) return rc;
  CatBest* bb = reinterpret_cast<CatBest*>(base);
  unsigned int* tk = reinterpret_cast<unsigned int*>(base + kStaticTicketsAt);
  // the array keeps the last call's contents: the tickets must start at zero (the winner slots are written before
  // they are read).  Stream-ordered, so the rollout below sees them cleared.
  CU(cudaMemsetAsync(tk, 0, (size_t)q.n_problems * sizeof(unsigned int), c->stream));
  CatNvtxRange r(c, "cat_mpc_step_static_dev");
  return mpc_launch<double>(c, q, problems, controls, p_ref, best_cost, (long long*)best_idx, costs, bb, tk);
}
