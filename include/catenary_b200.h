/* catenary_b200.h -- C ABI of libcatenary_b200.so (sm_100a).
 *
 * The reference (eather0056/Catenary-Model-Estimation-and-MPC-Control-for-ROV-Tethered-Systems)
 * is pure Python and has no FFI: its boundary for this path is a set of module-level functions
 * and one callable class.  Each entry point below names the reference interface it replaces
 * (file:line relative to the reference checkout).  INTEGRATION.md shows the ctypes stub a
 * maintainer of the reference would add.
 *
 * Conventions
 *  - return 0 (CAT_OK) on success, a negative CAT_E* code otherwise; never throws, never aborts.
 *    cat_last_error(ctx) gives a message valid until the next call on that ctx.
 *  - numerical failure is DATA, as in the reference: solve -> NaN (main_fun.py:427-428),
 *    shape outside the solve domain -> valid[i] = 0 (callers substitute [A,B], main_fun.py:67-69).
 *  - the caller owns every buffer.  Un-suffixed functions take HOST pointers and do the
 *    host<->device copies on the ctx stream; *_dev functions take DEVICE pointers, enqueue on the
 *    ctx stream and return without synchronising (cat_sync to wait).
 *  - one ctx = one device + one stream; calls on one ctx are serialised by the caller.
 *  - there is NO CPU fallback: without a CUDA device cat_init fails with CAT_ENODEV.
 *  - objects created on a ctx (cat_ctrl_create_*, cat_p2p_create) hold a pointer to it: destroy them before
 *    cat_destroy(ctx).
 */
#ifndef CATENARY_B200_H
#define CATENARY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CAT_ABI_VERSION 3

#if defined(CAT_BUILD_SHARED) && defined(__GNUC__)
#define CAT_API __attribute__((visibility("default")))
#else
#define CAT_API
#endif

enum {
  CAT_OK = 0,
  CAT_EINVAL = -1,   /* bad argument (null pointer, negative size, bad enum) */
  CAT_ECUDA = -2,    /* CUDA runtime error (message in cat_last_error) */
  CAT_ENOMEM = -3,   /* device or pinned allocation failed */
  CAT_ENODEV = -4    /* no usable CUDA device */
};

/* which outputs of transform_catenary (main_fun.py:111) to write, in this order */
enum {
  CAT_OUT_ORIGINAL = 1,
  CAT_OUT_THETA_ROTATED = 2,
  CAT_OUT_THETA_ALIGNED = 4,
  CAT_OUT_FINAL = 8
};

/* flags of CatMpcParams */
enum {
  CAT_MPC_RNG_CONTROLS = 1,  /* controls == NULL: sample on device (Philox4x32-10 + tabulated inverse normal CDF) */
  CAT_MPC_CONTROLS_AOS = 2,  /* controls are [n_problems][K][H][3] (the oracle's (K,H,3)) instead of SoA */
  CAT_MPC_CONTROLS_F32 = 4   /* *_f64 entry points only: the controls buffer holds float, not double (the pointer
                                argument is reinterpreted).  Arithmetic stays fp64; the candidates cross PCIe and HBM
                                at half the bytes.  A float rollout (*_f32) always reads float controls. */
};

/* modes of the fused multi-GPU winner exchange (cat_p2p_set_mode) */
enum {
  CAT_P2P_MODE_LOCKSTEP = 0, /* a step returns the job-wide winner of ITSELF: its kernel ends when every rank's
                                winner has arrived (the latency path) */
  CAT_P2P_MODE_PIPELINED = 1 /* a step returns the job-wide winner of the PREVIOUS step and never waits for a peer
                                that is less than a whole step behind (the throughput path); cat_p2p_flush_dev
                                delivers the winner of the last step */
};

/* receding-horizon controller: how the next nominal sequence is formed */
enum {
  CAT_CTRL_ARGMIN = 0,       /* the best candidate of this step */
  CAT_CTRL_MPPI = 1          /* exp(-(J - Jmin)/lambda)-weighted mean of all candidates */
};
#define CAT_CTRL_LOG_W 16    /* doubles per logged control step, see cat_ctrl_read */

/* Shared parameters of one sampling-MPC step (builder-defined composition).
 * Plain-old-data, passed by pointer from the host, by value to the kernel. */
typedef struct CatMpcParams {
  int64_t K;            /* candidates per problem evaluated by THIS call (the local shard) */
  int64_t k_offset;     /* global index of local candidate 0 (multi-GPU shards; RNG counter, best_idx) */
  int32_t H;            /* horizon steps */
  int32_t n_problems;   /* >= 1; fleet mode when > 1 */
  int32_t flags;        /* CAT_MPC_* */
  int32_t problem_offset; /* global index of local problem 0 (fleet shards; RNG counter) */
  int32_t eq_theta;     /* generated equation id of d(theta)/dt, -1 = shipped (eq_dtheta_dt.txt) */
  int32_t eq_gamma;     /* generated equation id of d(gamma)/dt, -1 = shipped (eq_dgamma_dt.txt) */
  double dt;            /* s */
  double L;             /* cable length, m (catenary.py:10) */
  double w_trk, w_du, w_len, w_sag, w_ang;
  double rho, z_floor, margin;
  double c_lo, c_hi;    /* bracket of main_fun.py:425: C outside (c_lo,c_hi) is "nan" */
  double nan_penalty;
  uint64_t seed;        /* RNG key */
  double sigma[3];      /* mm/s, per axis */
  double clip;          /* in sigmas */
} CatMpcParams;

/* Per-problem initial state (fully_augmented_catenary.py:11-15 for the nominal pose). */
typedef struct CatMpcProblem {
  double P0[3];         /* anchor, m */
  double P1[3];         /* ROV attach point, m */
  double V1[3];         /* mm/s (simple.py:18 leaves velocities in data units) */
  double A1[3];         /* mm/s^2 */
  double theta, gamma, theta_prev, gamma_prev;   /* rad */
} CatMpcProblem;

/* ---- context ------------------------------------------------------------------------- */
CAT_API int cat_abi_version(void);
CAT_API int cat_init(int device, void** ctx);
CAT_API int cat_destroy(void* ctx);
CAT_API int cat_sync(void* ctx);
CAT_API const char* cat_last_error(void* ctx);
/* number of kernels this ctx has launched so far (bench.py's gpu_launches) */
CAT_API int64_t cat_launch_count(void* ctx);
/* the ctx stream (cudaStream_t) for callers that enqueue their own work around *_dev calls */
CAT_API void* cat_stream(void* ctx);
/* run subsequent calls of this ctx on a caller-owned stream; own != 0 restores the ctx stream (cuda_stream is
 * then ignored).  cuda_stream == NULL with own == 0 is the CUDA legacy default stream. */
CAT_API int cat_set_stream(void* ctx, void* cuda_stream, int own);
/* pinned host memory for the un-suffixed (host-pointer) entry points */
CAT_API int cat_host_alloc(void* ctx, size_t bytes, void** out);
CAT_API int cat_host_free(void* ctx, void* p);
/* page-lock a CALLER-owned host range (an ordinary NumPy array) so that the host-pointer entry points copy it at
 * PCIe speed instead of through the driver's pageable staging path; the caller unregisters before freeing the
 * memory (the Python layer ties it to the array's lifetime).  At most 64 ranges per ctx. */
CAT_API int cat_host_register(void* ctx, void* p, size_t bytes);
CAT_API int cat_host_unregister(void* ctx, void* p);
/* Private per-block winner slots + tickets for the *_dev MPC entry points.  The default (ws == NULL) is scratch
 * shared by every call on the ctx, which is only safe when launches of one ctx never overlap; a caller that
 * pipelines steps over several streams, or captures a step into a CUDA graph, creates one workspace per in-flight
 * step and selects it with cat_set_workspace before the launch (like cat_set_stream).  Sized for max_candidates x
 * max_problems of one launch. */
CAT_API int cat_mpc_workspace_create(void* ctx, int64_t max_candidates, int32_t max_problems, void** ws);
CAT_API int cat_mpc_workspace_destroy(void* ctx, void* ws);
CAT_API int cat_set_workspace(void* ctx, void* ws_or_null);
/* Rollout kernels are launched with programmatic stream serialization (default on; env CAT_MPC_OVERLAP=0 or on = 0
 * turns it off): back-to-back steps on one stream overlap -- the next step's candidates start on the SMs the previous
 * step's early finishers leave idle and run under its reduction / exchange tail.  Stream order is kept for everything a
 * caller can observe: a step touches its outputs and the shared scratch only after the preceding step has completed,
 * and a step that follows any other kind of work on the stream (a copy, another library's kernel) starts after it. */
CAT_API int cat_set_overlap(void* ctx, int on);
/* cold-path counters of the rollout kernel (bench.py "sensitivity"): enable != 0 switches counting on, 0 off;
 * out (may be NULL) receives, summed over every rollout launched by this process since the last reset:
 * [0] warp-level entries into the catenary solve's iteration loop (the warm Halley step did not suffice),
 * [1] warp-level entries into the full-range sine (|x| >= 1e5, inf, NaN), [2], [3] reserved.  reset != 0 zeroes
 * them after reading.  Synchronises the device. */
CAT_API int cat_mpc_stats(void* ctx, int enable, int reset, uint64_t* out);
/* device event timing on the ctx stream (bench.py; torch.cuda.Event only sees torch's stream) */
CAT_API int cat_timer_start(void* ctx);
CAT_API int cat_timer_stop_ms(void* ctx, float* ms);   /* records, synchronises, returns elapsed */

/* ---- catenary parameter: replaces main_fun.solve_catenary (main_fun.py:418-431) -------
 * C_out[i] = root C in (1e-6, 10) of C^2 (L^2 - dH^2) - 4 sinh^2(l C / 2) = 0, NaN where the
 * reference's brentq bracket has no sign change.  L_stride = 0 broadcasts L[0]. */
CAT_API int cat_solve_param_f64(void* ctx, int64_t n, const double* l, const double* dH,
                        const double* L, int64_t L_stride, double* C_out);
CAT_API int cat_solve_param_f32(void* ctx, int64_t n, const float* l, const float* dH,
                        const float* L, int64_t L_stride, float* C_out);
CAT_API int cat_solve_param_dev_f64(void* ctx, int64_t n, const double* l, const double* dH,
                            const double* L, int64_t L_stride, double* C_out);
CAT_API int cat_solve_param_dev_f32(void* ctx, int64_t n, const float* l, const float* dH,
                            const float* L, int64_t L_stride, float* C_out);

/* span terms in one pass: C as above; x_low = horizontal offset of the lowest point from A
 * (l/2 - atanh(dH/L)/C, NaN where C is NaN); z_low = lowest z relative to zA (min(0,dH) when the vertex is
 * outside the span or C is NaN); tension = w l / (2 sinh(C l / 2)), NaN C -> w l / 2 (main_fun.py:302-305,
 * w = cable wet weight, test_cluster.py:23).  Any output pointer may be NULL. */
CAT_API int cat_span_terms_f64(void* ctx, int64_t n, const double* l, const double* dH, const double* L,
                               int64_t L_stride, double w, double* C_out, double* x_low, double* z_low,
                               double* tension);
/* replaces main_fun.rodrigues_rotation (main_fun.py:18-35), batched: out[i] = origin + R(axis, angle)(v[i] - origin).
 * v, out (n,3); axis (n,3) or one axis (axis_stride 0); angle n values or one (angle_stride 0);
 * origin 3 values or NULL. */
CAT_API int cat_rodrigues_f64(void* ctx, int64_t n, const double* v, const double* axis, int64_t axis_stride,
                              const double* angle, int64_t angle_stride, const double* origin, double* out);

/* ---- curve shapes: replaces Catenary.__call__ (pympc, call sites catenary.py:25,51) +
 *      main_fun.transform_catenary (main_fun.py:38-111) + rodrigues_rotation (main_fun.py:18-35)
 * A, B: AoS (n,3).  L: n values or one (L_stride 0).  theta, gamma: n values, or NULL (= 0).
 * which: bitmask of CAT_OUT_*; out holds popcount(which) consecutive blocks, each AoS
 * (n, n_pts, 3), in the enum's order.  C_out (n) and valid (n) may be NULL.
 * valid bit 0: original span A->B is in the solve domain; bit 1: rotated span A->B' is.
 * A block whose span is outside the domain holds the straight chord sampled at n_pts points. */
CAT_API int cat_shape_f64(void* ctx, int64_t n, const double* A, const double* B, const double* L,
                  int64_t L_stride, const double* theta, const double* gamma, int32_t n_pts,
                  int32_t which, double* out, double* C_out, uint8_t* valid);
CAT_API int cat_shape_f32(void* ctx, int64_t n, const float* A, const float* B, const float* L,
                  int64_t L_stride, const float* theta, const float* gamma, int32_t n_pts,
                  int32_t which, float* out, float* C_out, uint8_t* valid);
CAT_API int cat_shape_dev_f64(void* ctx, int64_t n, const double* A, const double* B, const double* L,
                      int64_t L_stride, const double* theta, const double* gamma, int32_t n_pts,
                      int32_t which, double* out, double* C_out, uint8_t* valid);
CAT_API int cat_shape_dev_f32(void* ctx, int64_t n, const float* A, const float* B, const float* L,
                      int64_t L_stride, const float* theta, const float* gamma, int32_t n_pts,
                      int32_t which, float* out, float* C_out, uint8_t* valid);
/* lowest sample of each curve: replaces np.argmin(curve[:,2]) (fully_augmented_catenary.py:20-22).
 * curves AoS (n, n_pts, 3) on the host; idx_out (n) first-minimum index, pt_out (n,3). */
CAT_API int cat_lowest_point_f64(void* ctx, int64_t n, int32_t n_pts, const double* curves,
                         int64_t* idx_out, double* pt_out);

/* ---- equations + open-loop integrators ------------------------------------------------
 * eq_id indexes the generated hall-of-fame table (cat_equation_count/info); the shipped pair is
 * cat_equation_shipped(0) = dtheta_dt, cat_equation_shipped(1) = dgamma_dt. */
CAT_API int cat_equation_count(void);
CAT_API int cat_equation_shipped(int which);
CAT_API int cat_equation_info(int eq_id, int* is_gamma, int* complexity, uint32_t* live_mask);
/* replaces PySRRegressor.predict on (already scaled) features: X (n,18) row-major -> out (n) */
CAT_API int cat_predict_f64(void* ctx, int64_t n, const double* X, int eq_id, int apply_scaler, double* out);
CAT_API int cat_predict_f32(void* ctx, int64_t n, const float* X, int eq_id, int apply_scaler, float* out);
/* replaces rk4_integration (simulate_rk4_theta_gamma.py:52-68): X (n_traj,T,18), time (T),
 * y0 (n_traj) -> y_out (n_traj,T).  method 0 = RK4 (k2 == k3 as in the reference),
 * 1 = explicit Euler (main_fun.integrate_theta_gamma, main_fun.py:735-764). */
CAT_API int cat_rk4_open_f64(void* ctx, int64_t n_traj, int64_t T, const double* X, const double* time,
                     const double* y0, int eq_id, int method, double* y_out);
CAT_API int cat_rk4_open_f32(void* ctx, int64_t n_traj, int64_t T, const float* X, const float* time,
                     const float* y0, int eq_id, int method, float* y_out);
/* replaces extract_features (simple.py:15-41) on columns already in SI/data units:
 * P0,P1 (T,3) m; V1 (T,3) mm/s; time, theta, gamma (T) -> X (T,18); scaled if apply_scaler. */
CAT_API int cat_features_f64(void* ctx, int64_t T, const double* P0, const double* P1, const double* V1,
                     const double* time, const double* theta, const double* gamma,
                     int apply_scaler, double* X_out);

/* ---- fused closed-loop sampling-MPC step (new composition) --------
 * problems: n_problems structs.  controls: [n_problems][H][3][K] (SoA over candidates, mm/s) or
 * NULL with CAT_MPC_RNG_CONTROLS.  p_ref: [n_problems][H][3] (m).
 * best_cost / best_idx: n_problems values (best_idx is GLOBAL: k_offset + local index; lowest cost, ties -> lowest
 * index, a NaN cost ranks with +inf so that a failed rollout is never selected).  costs: [n_problems][K] or NULL. */
/* Host-pointer step with one problem and >= 32,768 explicit candidates: the upload overlaps the rollout.  SoA
 * candidates of the shipped equation pair whose grid is resident at once (K <= 4 * 128 * SMs = 75,776 on B200, rows
 * of K values a multiple of 128 bytes, <= 1 GB in all, in page-locked host memory) are STREAMED: one rollout is launched at once and reads each
 * horizon step's rows after their row group has landed (arrival words written with a stream memory operation);
 * everything else goes up in two candidate chunks (the last one block per SM) with one rollout each.  Identical
 * results either way.
 * Env CAT_MPC_STREAMED=0 (read in cat_init) forces the chunked form. */
CAT_API int cat_mpc_step_f64(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems,
                     const double* controls, const double* p_ref,
                     double* best_cost, int64_t* best_idx, double* costs);
CAT_API int cat_mpc_step_f32(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems,
                     const float* controls, const double* p_ref,
                     double* best_cost, int64_t* best_idx, float* costs);
/* device-pointer variants: problems, controls, p_ref, best_cost, best_idx, costs all on device */
CAT_API int cat_mpc_step_dev_f64(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems,
                         const double* controls, const double* p_ref,
                         double* best_cost, int64_t* best_idx, double* costs);
CAT_API int cat_mpc_step_dev_f32(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems,
                         const float* controls, const double* p_ref,
                         double* best_cost, int64_t* best_idx, float* costs);
/* Allocation-free variant of cat_mpc_step_dev_f64 (same arguments, same results).  It has NO counterpart in the
 * reference: it was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The rollout's scratch (one winner slot per
 * thread block, one ticket per problem) lives in a fixed file-scope device array of 528 KB that every GPU loading the
 * library holds until exit; the call makes no cudaMalloc / cudaFree and ignores cat_set_workspace.
 * Limits: ceil(K / 128) * n_problems <= CAT_STATIC_MAX_BLOCKS and n_problems <= CAT_STATIC_MAX_PROBLEMS, else
 * CAT_EINVAL with a message in cat_last_error (run-time equation programs are not offered here).
 * All calls on one device share the array: calls that overlap on the GPU (different streams or contexts) race on it
 * and must be ordered by the caller; back-to-back calls on one stream are safe. */
#define CAT_STATIC_MAX_BLOCKS 32768
#define CAT_STATIC_MAX_PROBLEMS 4096
CAT_API int cat_mpc_step_static_dev_f64(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems,
                                const double* controls, const double* p_ref,
                                double* best_cost, int64_t* best_idx, double* costs);
/* materialise the on-device RNG controls ([n_problems][H][3][K], device pointer) */
CAT_API int cat_sample_controls_dev_f64(void* ctx, const CatMpcParams* p, double* controls);
/* deterministic reduce of per-shard winners gathered from G ranks (cost[g], idx[g]):
 * lowest cost, ties -> lowest global index, NaN ranks with +inf.  Host arrays. */
CAT_API int cat_reduce_winners_f64(int n, const double* cost, const int64_t* idx,
                           double* best_cost, int64_t* best_idx);

/* the same fold on the device (enqueued on the ctx stream, no host sync): gathered = n x [cost bits, global idx]
 * as int64 pairs (what cat_mpc_step_dev_* leaves when best_cost and best_idx point into one 16-byte buffer and
 * the ranks all-gather it), out = one such pair. */
CAT_API int cat_reduce_winners_dev_f64(void* ctx, int n, const int64_t* gathered, int64_t* out);

/* replaces compute_catenary_3D (models/catenary_3d.py:5-39), batched: p0, p1 (n,3), rope length L (n or one,
 * L_stride 0) -> out (n, n_pts, 3).  Taut rope (L <= |p1-p0|) -> np.linspace(p0, p1, n_pts); else the reference's
 * fixed point a <- a L / (2 a sinh(d/2a)) (<= 100 passes, 1e-6) and its symmetric sag profile. */
CAT_API int cat_catenary3d_f64(void* ctx, int64_t n, const double* p0, const double* p1, const double* L,
                               int64_t L_stride, int32_t n_pts, double* out);
/* replaces catenary_equation (models/catenary-3d-visualization/src/catenary_model.py:10-12):
 * y[i] = (cosh(C x[i]) - 1) / C, C one value (C_stride 0) or n. */
CAT_API int cat_catenary_equation_f64(void* ctx, int64_t n, const double* x, const double* C, int64_t C_stride,
                                      double* y);

/* replaces uniform_resample (main_fun.py:128-145): scipy interp1d(kind='linear', fill_value='extrapolate') of every
 * column onto a new time grid.  time_in (n_in, increasing), cols (n_in, n_cols) row-major like DataFrame.values,
 * time_out (n_out) -> out (n_out, n_cols); bit-identical to scipy 1.18's interp1d._call_linear. */
CAT_API int cat_resample_linear_f64(void* ctx, int64_t n_in, const double* time_in, int32_t n_cols,
                                    const double* cols, int64_t n_out, const double* time_out, double* out);

/* feature term bank: every per-row term of the reference's other feature builders --
 * build_theta_features / build_gamma_features (main_fun.py:267-415), build_clean_features (main_fun.py:197-265),
 * extract_features (main_fun.py:166-193) -- in one pass, including the per-row solve_catenary + tension that the
 * reference runs through np.vectorize(brentq).  Inputs as cat_features_f64; L cable length (m), w cable wet
 * weight (N, test_cluster.py:23).  bank (T, 32) row-major:
 *  0-2 P1 | 3-5 V1 | 6-8 A1 = np.gradient(V1, time) | 9-11 unit_rel | 12 clip(|rel|, 1e-5, 10) | 13 angle_proj |
 *  14 v_surge = V1.unit_rel | 15 v_sway = |V1 x unit_rel| | 16 l | 17 dH | 18 C | 19 T = (w/L) l / (2 sinh(C l / 2)),
 *  nan C -> (w/L) l / 2 | 20 tanh(angle_proj) | 21 A1.unit_rel | 22 |A1 - (A1.unit_rel) unit_rel| | 23 |V1| | 24 |A1| |
 *  25 |A1| / (|V1| + 1e-8) | 26 theta | 27 gamma | 28 theta_prev | 29 gamma_prev | 30 |rel| | 31 0 */
CAT_API int cat_feature_bank_f64(void* ctx, int64_t T, const double* P0, const double* P1, const double* V1,
                                 const double* time, const double* theta, const double* gamma, double L, double w,
                                 double* bank);

/* ---- smoothed / second-order feature layouts (second half) -------------------
 * Signals are stored one after the other: x, y, out are [n_sig][len].  The filter constants come from the caller
 * (features.py computes them like SciPy does), so sigma, window length and polynomial order are free.
 * cat_fir_reflect: replaces scipy.ndimage.gaussian_filter1d as the reference calls it (main_fun.py:510-519, 772-773):
 *   correlation with n_taps (odd) symmetric taps, boundary mode 'reflect' (d c b a | a b c d | d c b a), any len >= 1.
 * cat_savgol: replaces scipy.signal.savgol_filter(x, window, order), mode='interp' (main_fun.py:649-650, 829-830):
 *   interior = taps (window) . x[i-h..i+h]; first / last h = window/2 samples = edge_l / edge_r (h x window, row-major)
 *   applied to the first / last `window` samples (the window's least-squares polynomial evaluated in place).
 * cat_gradient: replaces np.gradient(y, time) (edge_order 1): uniform != 0 selects NumPy's uniform branch
 *   ((f[i+1] - f[i-1]) / 2h, taken when every spacing of `time` is bitwise equal), else the second-order
 *   non-uniform formula; len >= 2. */
CAT_API int cat_fir_reflect_f64(void* ctx, int64_t n_sig, int64_t len, const double* x, int32_t n_taps,
                                const double* taps, double* out);
CAT_API int cat_savgol_f64(void* ctx, int64_t n_sig, int64_t len, const double* x, int32_t window, const double* taps,
                           const double* edge_l, const double* edge_r, double* out);
CAT_API int cat_gradient_f64(void* ctx, int64_t n_sig, int64_t len, const double* y, const double* time,
                             int32_t uniform, double* out);
/* replaces build_theta_features_valid (for_gamma == 0, main_fun.py:504-571) / build_gamma_features_valid
 * (for_gamma != 0, main_fun.py:574-642): inputs as cat_features_f64, taps = the Gaussian (sigma = 2 in the reference)
 * -> out (T,13) row-major.  One upload, five kernels, one download. */
CAT_API int cat_features_valid_f64(void* ctx, int64_t T, const double* P0, const double* P1, const double* V1,
                                   const double* time, const double* theta, const double* gamma, int32_t n_taps,
                                   const double* taps, int32_t uniform, int32_t for_gamma, double* out);
/* replaces features_dd (main_fun.py:811-871): Savitzky-Golay smoothed angles, their first and second np.gradient,
 * cable-relative surge / sway velocities (V1 / 1000: m/s) and their gradients -> feats (T,14), targets (T,2). */
CAT_API int cat_features_dd_f64(void* ctx, int64_t T, const double* P0, const double* P1, const double* V1,
                                const double* time, const double* theta, const double* gamma, int32_t window,
                                const double* taps, const double* edge_l, const double* edge_r, int32_t uniform,
                                double* feats, double* targets);

/* ---- run-time equation programs ----------------------------------
 * Any PySR hall-of-fame row -- the ~3,900 equations under the reference's cluster_run/outputs and outputs
 * directories, or a model the user retrains (simple.py:60-75) -- evaluated without rebuilding the library:
 * the host compiles the CSV's sympy_format / Equation string to a postfix program (expr.py) and every GPU
 * thread runs it on its feature row.  ops: n_ops pairs (opcode, arg), opcodes:
 *   0 CONST(arg = index into consts)  1 VAR(arg = feature index)  2 ADD 3 SUB 4 MUL 5 DIV 6 NEG 7 SIN 8 COS
 *   9 TANH 10 EXP 11 LOG 12 SQRT 13 ABS 14 SQUARE 15 CUBE 16 SIGN 17 POW (a**b, both on the stack)
 *   18 POWI (a**arg, |arg| <= 64).  A malformed program (stack under/overflow, depth > 32, bad index) is
 *   CAT_EINVAL.
 * cat_expr_eval: replaces PySRRegressor.predict (test_temp.py:53-54): X (n, n_features) row-major -> out (n);
 *   mean/scale (n_features each, both or neither): StandardScaler.transform applied first (test_temp.py:49).
 * cat_expr_integrate: replaces rk4_integration (simulate_rk4_theta_gamma.py:52-68, method 0),
 *   integrate_theta_gamma (main_fun.py:735-764, method 1) and the second-order scheme of
 *   test_cluster.py:110-129 / dd_test_cluster.py:80-99 (method 2: the equation predicts the second derivative;
 *   ydot starts at 0 and is Euler-integrated, then y from y0; ydot_out (n_traj,T) may be NULL).
 *   X (n_traj, T, n_features), time (T), y0 (n_traj) -> y_out (n_traj, T). */
CAT_API int cat_expr_eval_f64(void* ctx, int64_t n, int32_t n_features, const double* X, int32_t n_ops,
                              const int32_t* ops, int32_t n_consts, const double* consts, const double* mean,
                              const double* scale, double* out);
CAT_API int cat_expr_eval_f32(void* ctx, int64_t n, int32_t n_features, const float* X, int32_t n_ops,
                              const int32_t* ops, int32_t n_consts, const double* consts, const double* mean,
                              const double* scale, float* out);
CAT_API int cat_expr_integrate_f64(void* ctx, int64_t n_traj, int64_t T, int32_t n_features, const double* X,
                                   const double* time, const double* y0, int32_t n_ops, const int32_t* ops,
                                   int32_t n_consts, const double* consts, int32_t method, double* y_out,
                                   double* ydot_out);

/* the sampling-MPC step with RUN-TIME equation programs for d(theta)/dt and d(gamma)/dt over the 18-feature layout
 * of simple.py:15-41 (a model retrained with simple.py:60-75 is usable without rebuilding the library); p->eq_theta /
 * p->eq_gamma are ignored.  Programs as in cat_expr_eval (n_features = 18).  Everything else as cat_mpc_step_*. */
CAT_API int cat_mpc_step_expr_f64(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems,
                                  const double* controls, const double* p_ref, int32_t n_ops_theta,
                                  const int32_t* ops_theta, int32_t n_consts_theta, const double* consts_theta,
                                  int32_t n_ops_gamma, const int32_t* ops_gamma, int32_t n_consts_gamma,
                                  const double* consts_gamma, double* best_cost, int64_t* best_idx, double* costs);
CAT_API int cat_mpc_step_expr_f32(void* ctx, const CatMpcParams* p, const CatMpcProblem* problems,
                                  const float* controls, const double* p_ref, int32_t n_ops_theta,
                                  const int32_t* ops_theta, int32_t n_consts_theta, const double* consts_theta,
                                  int32_t n_ops_gamma, const int32_t* ops_gamma, int32_t n_consts_gamma,
                                  const double* consts_gamma, double* best_cost, int64_t* best_idx, float* costs);

/* ---- fused multi-GPU winner exchange over NVLink peer memory --------------------
 * One process per GPU on one node.  Each rank creates its slot buffer (cat_p2p_create -> a 64-byte CUDA IPC
 * handle), the ranks exchange the handles by any means (torch.distributed all_gather) and map each other's
 * buffers (cat_p2p_connect, handles in rank order).  cat_mpc_step_p2p_dev_* is then ONE kernel per step: the
 * rollout's last block writes this rank's (cost, global idx) into every peer's buffer, waits for the peers'
 * winners of the same step and folds them in rank order, so best_cost / best_idx (device pointers, one value
 * each) hold the JOB-WIDE winner on every rank when the kernel ends -- no NCCL call, no second launch.  Every
 * rank must launch the same sequence of steps; n_problems must be 1.  A peer that does not show up within a few
 * seconds makes the step return the sentinel best_cost = NaN, best_idx = -1 (never a fold over stale slots). */
CAT_API int cat_p2p_create(void* ctx, int rank, int n_ranks, void** p2p, void* handle_out /* 64 bytes */);
CAT_API int cat_p2p_connect(void* p2p, const void* all_handles /* n_ranks x 64 bytes, rank order */);
CAT_API int cat_mpc_step_p2p_dev_f64(void* ctx, void* p2p, const CatMpcParams* p, const CatMpcProblem* problems,
                                     const double* controls, const double* p_ref, double* best_cost,
                                     int64_t* best_idx, double* costs);
CAT_API int cat_mpc_step_p2p_dev_f32(void* ctx, void* p2p, const CatMpcParams* p, const CatMpcProblem* problems,
                                     const float* controls, const double* p_ref, double* best_cost,
                                     int64_t* best_idx, float* costs);
/* host-pointer form of the same step (SURVEY.md 8b "cat_mpc_step_multi"): every rank passes its shard as host
 * arrays exactly as to cat_mpc_step_*, the library copies, rolls out, exchanges inside the kernel (lock-step) and
 * returns the JOB-WIDE winner in best_cost / best_idx on every rank.  n_problems must be 1. */
CAT_API int cat_mpc_step_p2p_f64(void* ctx, void* p2p, const CatMpcParams* p, const CatMpcProblem* problems,
                                 const double* controls, const double* p_ref, double* best_cost,
                                 int64_t* best_idx, double* costs);
CAT_API int cat_mpc_step_p2p_f32(void* ctx, void* p2p, const CatMpcParams* p, const CatMpcProblem* problems,
                                 const float* controls, const double* p_ref, double* best_cost,
                                 int64_t* best_idx, float* costs);
/* CAT_P2P_MODE_*; every rank must switch at the same step.  Not capturable into a CUDA graph (the step number is
 * a kernel argument). */
CAT_API int cat_p2p_set_mode(void* p2p, int mode);
/* job-wide winner of the most recent step (device pointers, one value each), without launching a rollout: the
 * closing call of a run of pipelined steps */
CAT_API int cat_p2p_flush_dev(void* ctx, void* p2p, double* best_cost, int64_t* best_idx);
/* synchronises, reports whether any step since the last call gave up waiting for a peer (such a step returns the
 * sentinel best_cost = NaN, best_idx = -1), and clears the flag */
CAT_API int cat_p2p_status(void* p2p, int* timed_out);
CAT_API int cat_p2p_destroy(void* p2p);

/* ---- receding-horizon controller around the sampling-MPC step ----
 * The role the (absent) pympc controller played at the reference's call sites: every control step samples K
 * candidate sequences around a nominal one, rolls them out (the kernel above), picks / blends the next nominal
 * sequence, applies its first control and shifts the rest by one step (warm start).  Everything -- plant state,
 * nominal sequence, step counter, log -- stays in HBM; a control step is 2 kernels (3 with MPPI), replayed as one
 * CUDA graph, with no host<->device traffic.  Definition: oracle mpc_receding_horizon.
 *   candidate k, step h:  u = clamp(u_nom[h] + sigma * clip(N(0,1)), +-u_max), Philox keyed by
 *   (seed + t, problem, k, h); candidate 0 carries no noise (the shifted previous solution itself).
 *   reference window of control step t: p_ref_traj[min(t + h, ref_len - 1)], h = 0..H-1.
 * p->flags is ignored (controls are always sampled on the device); p->k_offset must be 0. */
typedef struct CatCtrlConfig {
  int32_t mode;        /* CAT_CTRL_* */
  int32_t max_steps;   /* log capacity = most control steps this controller can run */
  int32_t use_graph;   /* != 0: capture one control step into a CUDA graph and replay it */
  int32_t reserved;
  double lambda;       /* MPPI temperature, cost units (> 0 when mode == CAT_CTRL_MPPI) */
  double u_max[3];     /* per-axis bound on the commanded velocity, mm/s */
} CatCtrlConfig;

/* problems: n_problems initial states (host).  u_nom0: [n_problems][H][3] host or NULL (= 0).
 * p_ref_traj: [n_problems][ref_len][3] host (m). */
CAT_API int cat_ctrl_create_f64(void* ctx, const CatMpcParams* p, const CatCtrlConfig* cfg,
                                const CatMpcProblem* problems, const double* u_nom0, const double* p_ref_traj,
                                int64_t ref_len, void** ctrl);
CAT_API int cat_ctrl_create_f32(void* ctx, const CatMpcParams* p, const CatCtrlConfig* cfg,
                                const CatMpcProblem* problems, const double* u_nom0, const double* p_ref_traj,
                                int64_t ref_len, void** ctrl);   /* fp32 rollouts, fp64 plant */
/* enqueue n_steps closed-loop control steps on the ctx stream (plant = the model itself); no host sync */
CAT_API int cat_ctrl_run(void* ctrl, int32_t n_steps);
/* overwrite the plant state with a measurement before the next step (n_problems structs, host) */
CAT_API int cat_ctrl_set_state(void* ctrl, const CatMpcProblem* problems);
/* synchronise and read back; any pointer may be NULL.  step_out: control steps done.  problems_out: current
 * plant states.  u_nom_out: [n_problems][H][3] current nominal sequence.  log_out: [n_problems][log_n]
 * [CAT_CTRL_LOG_W] rows of control steps log_from .. log_from+log_n-1:
 *   0-2 P1 after the step (m), 3-5 applied control (mm/s), 6 theta, 7 gamma, 8 C, 9 z_low, 10 best cost of
 *   the step's candidates, 11 best candidate index, 12 cost of the applied step, 13 step index, 14 span |P1-P0|,
 *   15 reserved. */
CAT_API int cat_ctrl_read(void* ctrl, int32_t* step_out, CatMpcProblem* problems_out, double* u_nom_out,
                          double* log_out, int32_t log_from, int32_t log_n);
CAT_API int cat_ctrl_destroy(void* ctrl);

/* ---- peak microbenchmarks for the roofline denominators (bench.py) -------------------
 * kind 0 = DFMA, 1 = FFMA, 2 = MUFU.EX2; returns operations per second (FMA = 1 op). */
CAT_API int cat_measure_peak(void* ctx, int kind, double* ops_per_s);

#ifdef __cplusplus
}
#endif
#endif /* CATENARY_B200_H */
