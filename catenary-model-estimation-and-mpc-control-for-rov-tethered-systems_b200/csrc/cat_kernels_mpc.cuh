// cat_kernels_mpc.cuh -- fused closed-loop sampling-MPC step (sm_100a).
//
// One thread per candidate control sequence.  The whole horizon runs in registers:
//   kinematics -> 18 features (simple.py:26-41) -> StandardScaler (test_temp.py:49) ->
//   RK4 stages of the generated d(theta)/dt, d(gamma)/dt (simulate_rk4_theta_gamma.py:61-66) ->
//   catenary parameter from (l, dH, L) (main_fun.py:423, warm-started from the previous step) ->
//   lowest point -> per-step cost -> running sum.
// Then a (cost, index) argmin: warp shuffle -> shared memory -> one slot per block -> the last
// block of each problem (atomic ticket) folds the slots.  Ordering is np.argmin's
// (fully_augmented_catenary.py:21): first minimum, NaN wins.
//
// The composition itself (closed loop, cost, argmin) has no reference counterpart; its definition
// is oracle/catenary_oracle.py::mpc_rollout_costs and this file follows that function line by line.
#pragma once
#include "cat_math.cuh"
#include "gen/dyn_eqs.cuh"
#include "cat_kernels_expr.cuh"
#include "../../include/catenary_b200.h"

#define CAT_MPC_BLOCK 128   // maximum; the host launches 64 when the batch is under ~1 wave (better SM balance)

// Fused winner exchange over NVLink peer memory: every rank owns a slot buffer
// [CAT_P2P_RING steps][n_ranks] x {cost bits, global idx, seq, pad} that its peers map through CUDA IPC.  The last
// block of the rollout POSTS this rank's winner of step s into slot [s % RING][rank] of EVERY rank's buffer
// (payload, system fence, then the sequence number) and then folds, in rank order, the G winners of
//   * step s itself (CAT_P2P_LOCKSTEP): it waits until its own buffer holds all of them, so the job-wide
//     (cost, idx) is in best_cost/best_idx when the kernel ends -- one kernel, no NCCL call, no second launch; the
//     latency path.  Every kernel then ends when the slowest rank's winner has arrived.
//   * step s - 1 (CAT_P2P_PIPELINED): those winners were posted one whole rollout ago, so nothing spins and
//     back-to-back steps of G independent processes no longer pay each other's launch jitter; the job-wide winner
//     of step s is delivered by step s + 1 (or by k_p2p_fold after the last step) -- the throughput path.
// Ring of 4: a rank finishes step s+1 only after it has seen every peer's post of s (pipelined) or s+1
// (lock-step), and a peer posts s only after its own fold of s-2 ended with its step-(s-1) kernel -- so nobody can
// be more than two posts ahead of a reader, and slot s % 4 is never rewritten (by s + 4) while s is still unread.
struct CatP2PDev {
  long long* const* peers;   // device array [n_ranks]: slot buffer of each rank (own included); null = off
  int rank, n_ranks;
  unsigned int seq;          // step number, >= 1, identical on every rank
  int mode;                  // CAT_P2P_LOCKSTEP / CAT_P2P_PIPELINED
  int* err;                  // set to 1 if a peer's winner did not arrive in time
};
#define CAT_P2P_SLOT_W 4     // long longs per slot
#define CAT_P2P_RING 4
#define CAT_P2P_MAX_RANKS 64
#define CAT_P2P_LOCKSTEP 0
#define CAT_P2P_PIPELINED 1

// Kernel-side parameters in the compute type.  Derived constants are formed on the host in
// double exactly as the oracle forms them (dt/1000, dt/6, rho*L, z_floor+margin).
#define CAT_ARRIVE_MAX 8   // row groups of the streamed upload
struct CatArrivePlan {
  unsigned int epoch;            // value the arrival words take in this call
  int n_groups;
  int end[CAT_ARRIVE_MAX];       // group g holds the horizon steps [end[g - 1], end[g]); the last end is H
  const unsigned int* words;     // arrival words (device memory)
  unsigned int* err;             // set to 1 when a group did not arrive in time (the costs are then junk)
};
template <typename T> struct CatMpcDev {
  long long K, k_offset;
  long long sk, sh, sa, sp;   // control strides in elements: candidate, step, axis, problem
  int H, n_problems, flags, problem_offset;
  int eq_theta, eq_gamma;
  int carry_t, carry_g;   // equation reads neither column 14 nor 15: k1 of the next step == k4 of this one
  int deep_clip;          // clip > CAT_ICDF_SMEM_ZMAX: the sampler's tail lanes read the full table (set by the host)
  T dt, dt_mm, dt6, L, w_trk, w_du, w_len, w_sag, w_ang, rhoL, rhoL2, z_thr, c_lo, c_hi, pen, clip;
  T L2;   // L^2 (sag-term bound)
  T sigma[3];
  unsigned long long seed;
  // receding-horizon controller mode (cat_kernels_ctrl.cuh); all null / H in a plain step
  const double* u_nom;    // [n_problems][H][3]: candidates = clamp(u_nom + noise, +-u_max); candidate 0 noise-free
  const int* step;        // device-resident control-step counter t: RNG key = seed + t, reference window starts at t
  int ref_len;            // rows of p_ref per problem (H in a plain step)
  T u_max[3];
  // run-time equation programs over the 18-feature layout (cat_kernels_expr.cuh); null = generated code.
  // prog = [theta ops | gamma ops] (int2), consts = [theta consts | gamma consts] (double), device memory.
  const int2* prog;
  const double* prog_consts;
  int n_ops_t, n_ops_g, n_c_t, n_c_g;
  int ref_smem_off;   // byte offset of the staged reference window in the dynamic shared memory
  CatP2PDev p2p;
  // streamed upload (host step, STREAM kernels): the candidates arrive over PCIe WHILE the rollout runs.  The host
  // copies the [H][3][K] rows in groups of whole horizon steps and, after each group, writes the call's epoch into
  // words[group] on the same stream; the kernel reads the controls of step h only after it has seen that word.
  // The plan of the call sits in device memory: one pointer, LAST in the block, so the offsets every other rollout
  // reads stay what they were (a field in the middle cost the sampling kernel 6 registers and 1% of its time).
  const CatArrivePlan* arrive;
};

// shared-memory view of the staged programs (VM mode)
template <typename T> struct CatMpcVm {
  const int2* ops_t;
  const int2* ops_g;
  const T* c_t;
  const T* c_g;
  int n_t, n_g;
};

// shared-memory slot of the sampler table: the real thing in the sampling kernels, one byte elsewhere
template <typename T, bool RNG> struct CatIcdfSlot { typedef CatIcdfSmem<T> type; };
template <typename T> struct CatIcdfSlot<T, false> { typedef char type; };
// Asynchronous control prefetch (cp.async, global -> shared, no register in between): thread t owns column t of a
// ring s[slot][axis][t] and keeps CAT_MPC_PF horizon steps in flight.  A ring held in REGISTERS does not work: the
// moves that rotate it need the loaded data, so the warp waits on step h+3's load at step h+1 (round-2 ncu of the
// fp32 kernel: 72% of the long-scoreboard samples on exactly that move, 1.2 stall cycles per issued instruction).
#ifndef CAT_MPC_ASYNC_F64
#define CAT_MPC_ASYNC_F64 0   // 1: the fp64 kernels prefetch this way too (A/B switch; fp32 always does)
#endif
#define CAT_MPC_PF 4
template <typename S, bool ON> struct CatRing {
  S v[(CAT_MPC_PF + 1) * 3 * CAT_MPC_BLOCK];
};
template <typename S> struct CatRing<S, false> {
  char v[1];
};
template <typename S>
__device__ __forceinline__ void cat_cp_async(S* smem_dst, const S* gsrc) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
  if constexpr (sizeof(S) == 4)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gsrc) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cat_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cat_cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// candidate control of the receding-horizon controller: clamp(u_nom[h] + sigma*clip(N(0,1)), +-u_max);
// global candidate 0 carries no noise (oracle ctrl_candidates)
template <typename T, typename M>
__device__ __forceinline__ void cat_nominal_control(const CatMpcDev<T>& q, unsigned long long seed_eff, int prob,
                                                    long long kglob, int h, CatIcdfRef icdf, T clip, T out[3]) {
  cat_sample_control<T, M>(seed_eff, (uint32_t)(q.problem_offset + prob), (uint64_t)kglob, (uint32_t)h, q.sigma,
                           clip, q.deep_clip != 0, icdf, out);
  const double* un = q.u_nom + ((long long)prob * q.H + h) * 3;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const T z = kglob == 0 ? T(0) : out[a];
    const T v = T(un[a]) + z;
    out[a] = M::clamp_sym_(v, q.u_max[a]);
  }
}

template <typename T, bool GENERIC, bool VM = false>
__device__ __forceinline__ void cat_eval_pair(const CatMpcDev<T>& q, const CatMpcVm<T>& vm, const T (&x)[18], T& ft,
                                              T& fg) {
  typedef CatMath<T> M;
  if constexpr (VM) {
    CatExprRow<T> row;
    row.a = x;
    row.b = nullptr;
    row.mean = nullptr;
    row.inv_scale = nullptr;
    ft = cat_expr_run<T>(vm.ops_t, vm.n_t, vm.c_t, row);
    fg = cat_expr_run<T>(vm.ops_g, vm.n_g, vm.c_g, row);
  } else if constexpr (GENERIC) {
    ft = cat_eval_equation<T, M>(q.eq_theta, x);
    fg = cat_eval_equation<T, M>(q.eq_gamma, x);
  } else {
    ft = cat_dtheta_dt<T, M>(x);
    fg = cat_dgamma_dt<T, M>(x);
  }
}

// both equations on N feature rows at once (the stage-2/3 and stage-4 rows of one RK4 step)
template <typename T, bool GENERIC, int N, bool VM = false>
__device__ __forceinline__ void cat_eval_pair_v(const CatMpcDev<T>& q, const CatMpcVm<T>& vm, const T (&x)[N][18],
                                                T (&ft)[N], T (&fg)[N]) {
  typedef CatMath<T> M;
  if constexpr (VM) {
#pragma unroll
    for (int n = 0; n < N; ++n) cat_eval_pair<T, GENERIC, true>(q, vm, x[n], ft[n], fg[n]);
  } else if constexpr (GENERIC) {
    cat_eval_equation_v<T, M, N>(q.eq_theta, x, ft);
    cat_eval_equation_v<T, M, N>(q.eq_gamma, x, fg);
  } else {
    cat_dtheta_dt_v<T, M, N>(x, ft);
    cat_dgamma_dt_v<T, M, N>(x, fg);
  }
}

// StandardScaler of one feature as a single FMA (x * inv_scale - mean * inv_scale: within 1 ulp of the scaled
// magnitude of (x - mean) * inv_scale); fp64 reads the pair from constant memory
template <typename T>
__device__ __forceinline__ T cat_scale1(T v, int j) {
  if constexpr (sizeof(T) == 8) return fma(v, CAT_SCALER_FMA[2 * j], CAT_SCALER_FMA[2 * j + 1]);
  else return fmaf(v, CatScaler<T>::inv_scale(j), CatScaler<T>::offset(j));
}

template <typename T>
__device__ __forceinline__ void cat_scale18(T (&x)[18]) {
#pragma unroll
  for (int j = 0; j < 18; ++j) x[j] = cat_scale1<T>(x[j], j);
}

// block-wide (cost, idx) argmin in the MPC winner order (cat_better_mpc: NaN ranks with +inf); result valid in
// thread 0
__device__ __forceinline__ CatBest cat_block_argmin(double cost, long long idx) {
  __shared__ double s_cost[CAT_MPC_BLOCK / 32];
  __shared__ long long s_idx[CAT_MPC_BLOCK / 32];
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const double oc = __shfl_xor_sync(0xffffffffu, cost, off);
    const long long oi = __shfl_xor_sync(0xffffffffu, idx, off);
    if (cat_better_mpc(oc, oi, cost, idx)) {
      cost = oc;
      idx = oi;
    }
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) {
    s_cost[w] = cost;
    s_idx[w] = idx;
  }
  __syncthreads();
  if (w == 0) {
    const int nw = blockDim.x >> 5;
    cost = (lane < nw) ? s_cost[lane] : __longlong_as_double(0x7ff0000000000000LL);
    idx = (lane < nw) ? s_idx[lane] : 0x7fffffffffffffffLL;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const double oc = __shfl_xor_sync(0xffffffffu, cost, off);
      const long long oi = __shfl_xor_sync(0xffffffffu, idx, off);
      if (cat_better_mpc(oc, oi, cost, idx)) {
        cost = oc;
        idx = oi;
      }
    }
  }
  CatBest b;
  b.cost = cost;
  b.idx = idx;
  return b;
}

// Post this rank's winner of step p2p.seq to every rank, then fold the G winners of the step the mode names
// (see CatP2PDev).  Called by every thread of ONE block (>= 64 threads).  A peer that does not show up within
// ~2-4 s raises *p2p.err and the result is the sentinel (NaN, -1), never a fold over stale slots: every consumer
// sees the failure in the value itself.
__device__ __forceinline__ long long* cat_p2p_slot(long long* base, unsigned int seq, int n_ranks, int r) {
  return base + ((long long)(seq % CAT_P2P_RING) * n_ranks + r) * CAT_P2P_SLOT_W;
}

__device__ __forceinline__ void cat_p2p_fold(const CatP2PDev& p2p, unsigned int want, double* __restrict__ best_cost,
                                             long long* __restrict__ best_idx) {
  __shared__ long long s_x[2 * CAT_P2P_MAX_RANKS];
  __shared__ int s_bad;
  const int t = threadIdx.x;
  if (t == 0) s_bad = 0;
  __syncthreads();
  if (t < p2p.n_ranks && want >= 1u) {
    volatile long long* src = cat_p2p_slot(p2p.peers[p2p.rank], want, p2p.n_ranks, t);
    int spins = 0;
    while (src[2] != (long long)want) {
      __nanosleep(64);
      if (++spins > (1 << 25)) {       // ~2-4 s: a peer is missing
        *p2p.err = 1;
        s_bad = 1;
        break;
      }
    }
    __threadfence_system();
    s_x[2 * t] = src[0];
    s_x[2 * t + 1] = src[1];
  }
  __syncthreads();
  if (t == 0) {
    if (want < 1u || s_bad) {          // no such step (first pipelined launch) or a missing peer
      *best_cost = __longlong_as_double(0x7ff8000000000000LL);
      *best_idx = -1;
    } else {
      double fc = __longlong_as_double(s_x[0]);
      long long fi = s_x[1];
      for (int g = 1; g < p2p.n_ranks; ++g) {
        const double c = __longlong_as_double(s_x[2 * g]);
        const long long i = s_x[2 * g + 1];
        if (cat_better_mpc(c, i, fc, fi)) {
          fc = c;
          fi = i;
        }
      }
      *best_cost = fc;
      *best_idx = fi;
    }
  }
}

__device__ __forceinline__ void cat_p2p_finish(long long my_c, long long my_i, const CatP2PDev& p2p,
                                               double* __restrict__ best_cost, long long* __restrict__ best_idx) {
  const int t = threadIdx.x;
  if (t < p2p.n_ranks) {
    volatile long long* dst = cat_p2p_slot(p2p.peers[t], p2p.seq, p2p.n_ranks, p2p.rank);
    dst[0] = my_c;
    dst[1] = my_i;
    __threadfence_system();            // payload visible system-wide before the sequence number
    dst[2] = (long long)p2p.seq;
  }
  cat_p2p_fold(p2p, p2p.mode == CAT_P2P_PIPELINED ? p2p.seq - 1u : p2p.seq, best_cost, best_idx);
}

// grid-wide argmin of one problem: block fold -> one slot per block -> the last block (atomic ticket)
// folds the slots, writes the winner and resets the ticket (re-entrant, graph-replayable).
__device__ __forceinline__ void cat_grid_argmin(double jc, long long ji, int prob, CatBest* __restrict__ block_best,
                                                unsigned int* __restrict__ tickets, double* __restrict__ best_cost,
                                                long long* __restrict__ best_idx, const CatP2PDev& p2p) {
  const CatBest b = cat_block_argmin(jc, ji);
  __shared__ bool s_last;
  if (threadIdx.x == 0) {
    block_best[(long long)prob * gridDim.x + blockIdx.x] = b;
    __threadfence();
    const unsigned int t = atomicAdd(&tickets[prob], 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double bc = __longlong_as_double(0x7ff0000000000000LL);
  long long bi = 0x7fffffffffffffffLL;
  const volatile CatBest* slots = block_best + (long long)prob * gridDim.x;
  for (unsigned int s = threadIdx.x; s < gridDim.x; s += blockDim.x) {
    const double oc = slots[s].cost;
    const long long oi = slots[s].idx;
    if (cat_better_mpc(oc, oi, bc, bi)) {
      bc = oc;
      bi = oi;
    }
  }
  __syncthreads();   // s_cost/s_idx of cat_block_argmin are reused
  const CatBest r = cat_block_argmin(bc, bi);
  if (p2p.peers == nullptr) {
    if (threadIdx.x == 0) {
      best_cost[prob] = r.cost;
      best_idx[prob] = r.idx;
      tickets[prob] = 0u;   // ready for the next launch / graph replay
    }
    return;
  }
  // ---- fused exchange of the rank winners through peer memory (one problem per launch) ----
  __shared__ long long s_mine[2];
  if (threadIdx.x == 0) {
    s_mine[0] = __double_as_longlong(r.cost);
    s_mine[1] = r.idx;
    tickets[prob] = 0u;
  }
  __syncthreads();
  cat_p2p_finish(s_mine[0], s_mine[1], p2p, best_cost + prob, best_idx + prob);
}

// ---- streamed upload: wait for the row group that holds the controls of a horizon step ----
struct CatArrive {
  int g, end;   // last group waited for; first step NOT covered by the groups waited for so far
};
// thread 0 of the block polls the group's word in L2 (sleeping between polls: 512 blocks hammering one address would
// hot-spot its L2 slice under the copy engine's writes), the block follows through a barrier.  Bounded: after ~2 s
// the block raises *err of the plan and carries on with whatever is in memory; the host reports the failure.
template <typename T>
__device__ __forceinline__ void cat_arrive_wait(const CatMpcDev<T>& q, int g) {
  if (threadIdx.x == 0) {
    const unsigned int* f = q.arrive->words + g;
    const unsigned int epoch = q.arrive->epoch;
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
    if (v != epoch) {
      unsigned long long t0, t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
      for (;;) {
        __nanosleep(400);
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
        if (v == epoch) break;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > 2000000000ull) {
          atomicExch(q.arrive->err, 1u);
          break;
        }
      }
    }
  }
  __syncthreads();
}
// the controls of step hh are about to be requested (hh never decreases): block-uniform
template <typename T>
__device__ __forceinline__ void cat_arrive_need(const CatMpcDev<T>& q, CatArrive& st, int hh) {
  while (hh >= st.end) {
    ++st.g;
    cat_arrive_wait<T>(q, st.g);
    st.end = q.arrive->end[st.g];
  }
}
// a control value of the streamed upload: L2 only (the line was written by the copy engine during this kernel)
template <typename S> __device__ __forceinline__ S cat_ld_streamed(const S* p) { return __ldcg(p); }

// T = compute type, S = storage type of the explicit controls (S = float with T = double: the candidates cross
// PCIe and HBM at 12 B per rollout-step instead of 24 and are widened on load -- the oracle is fed the same
// rounded values, so parity is untouched).  RNG = true draws the controls on the device instead (S unused).
template <typename T, typename S, bool GENERIC, bool VM = false, bool RNG = false, bool STREAM = false>
#ifndef CAT_MPC_MINB
#define CAT_MPC_MINB 4   // resident blocks per SM the register budget is cut for (A/B switch)
#endif
__global__ void __launch_bounds__(CAT_MPC_BLOCK, CAT_MPC_MINB)   // <= 128 registers: config 3 needs 3.5 blocks per SM resident
k_mpc_rollout(const CatMpcDev<T> q, const CatMpcProblem* __restrict__ problems, const S* __restrict__ controls,
              const double* __restrict__ p_ref, T* __restrict__ costs, CatBest* __restrict__ block_best,
              unsigned int* __restrict__ tickets, double* __restrict__ best_cost,
              long long* __restrict__ best_idx) {
  typedef CatMath<T> M;
  // Programmatic dependent launch (the host launches with cudaLaunchAttributeProgrammaticStreamSerialization): as
  // soon as every block of this grid is resident, the NEXT rollout on the stream may start on whatever the grid
  // leaves free -- 65,536 candidates are 512 blocks on 592 slots, and the 3-block SMs finish ~20% before the 4-block
  // ones -- and it runs under this grid's reduction / exchange tail.  Inputs of consecutive steps are independent;
  // everything shared is guarded by the griddepcontrol.wait behind the horizon loop.  No-ops without the attribute.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // inverse-CDF table of the device sampler (hot octaves), staged once per block
  __shared__ typename CatIcdfSlot<T, RNG>::type s_icdf;
  CatIcdfRef icdf = 0;
  if constexpr (RNG) {
    cat_icdf_stage<T>(s_icdf);
    icdf = cat_pin(cat_icdf_ref_of<T>(s_icdf));
  }
  const T clip_r = RNG ? cat_pin(q.clip) : q.clip;
  constexpr bool ASYNC = !RNG && (sizeof(T) == 4 || CAT_MPC_ASYNC_F64 != 0);
  __shared__ __align__(16) CatRing<S, ASYNC> s_ring;
  // dynamic shared memory: [theta ops | gamma ops | constants] of the run-time programs (VM), then the reference window
  extern __shared__ __align__(16) unsigned char cat_mpc_smem[];
  CatMpcVm<T> vm;
  if constexpr (VM) {   // stage both programs into shared memory once per block
    int2* s_ops = reinterpret_cast<int2*>(cat_mpc_smem);
    T* s_c = reinterpret_cast<T*>(s_ops + q.n_ops_t + q.n_ops_g);
    for (int i = threadIdx.x; i < q.n_ops_t + q.n_ops_g; i += blockDim.x) s_ops[i] = q.prog[i];
    for (int i = threadIdx.x; i < q.n_c_t + q.n_c_g; i += blockDim.x) s_c[i] = T(q.prog_consts[i]);
    vm.ops_t = s_ops;
    vm.ops_g = s_ops + q.n_ops_t;
    vm.c_t = s_c;
    vm.c_g = s_c + q.n_c_t;
    vm.n_t = q.n_ops_t;
    vm.n_g = q.n_ops_g;
  } else {
    vm.ops_t = vm.ops_g = nullptr;
    vm.c_t = vm.c_g = nullptr;
    vm.n_t = vm.n_g = 0;
  }
  const int prob = blockIdx.y;
  const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = k < q.K;
  const long long kk = live ? k : q.K - 1;   // idle lanes shadow the last candidate (no divergence, no OOB)

  const CatMpcProblem* pb = problems + prob;
  const T P0[3] = {T(pb->P0[0]), T(pb->P0[1]), T(pb->P0[2])};
  T P[3] = {T(pb->P1[0]), T(pb->P1[1]), T(pb->P1[2])};
  T V[3] = {T(pb->V1[0]), T(pb->V1[1]), T(pb->V1[2])};
  T A[3] = {T(pb->A1[0]), T(pb->A1[1]), T(pb->A1[2])};
  T th = T(pb->theta), ga = T(pb->gamma), thp = T(pb->theta_prev), gap = T(pb->gamma_prev);
  T J = T(0);
  // warm-start record of the catenary solve, seeded from the initial pose (a step moves the span by ~1e-3)
  CatWarm<T> warm;
  {
    const T wx = P[0] - P0[0], wy = P[1] - P0[1];
    warm = cat_warm_guess<T, M>(M::sqrt_(wx * wx + wy * wy), P[2] - P0[2], q.L);
  }

  // control source fixed at compile time: the explicit-controls kernel carries no Philox / sampler code and the
  // sampling kernel no control loads
  constexpr bool rng = RNG;
  const S* uc = rng ? nullptr : controls + (long long)prob * q.sp + kk * q.sk;
  // controller mode: the step counter lives in HBM so that one captured graph serves every control step
  int t0 = 0;
  if (q.step) t0 = min(*q.step, q.ref_len - 1);
  unsigned long long seed_eff = q.seed + (unsigned long long)(q.step ? *q.step : 0);
  if constexpr (RNG) {
    // the key is the same in every lane, but a value loaded from HBM lives in a vector register: the ten Philox round
    // keys were then re-derived with one integer add per round and lane.  A warp reduction returns it in a uniform
    // register, and the round keys fold into the uniform operand of the xor.  (All lanes are here: no lane has exited.)
    const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)seed_eff);
    const unsigned hi = __reduce_or_sync(0xffffffffu, (unsigned)(seed_eff >> 32));
    seed_eff = ((unsigned long long)hi << 32) | lo;
  }
  const int ref_rem = q.ref_len - t0;   // rows left in the reference trajectory; the last one is held
  const double* pr = p_ref + ((long long)prob * q.ref_len + t0) * 3;
  // The H reference points of this launch, in the compute type, staged once per block: the loop reads them with a
  // broadcast LDS.  (Fetched from global memory one step ahead into loop-carried registers, the move that rotates the
  // register waits for the load: round-2 ncu of the fp32 kernel, 60% of its long-scoreboard samples.)
  T* const s_ref = reinterpret_cast<T*>(cat_mpc_smem + q.ref_smem_off);
  for (int i = threadIdx.x; i < 3 * q.H; i += blockDim.x) {
    const int hh = i / 3;
    const int row = hh < ref_rem ? hh : ref_rem - 1;
    s_ref[i] = T(pr[3 * row + (i - 3 * hh)]);
  }
  __syncthreads();   // programs (VM) and reference window are in place

  // Scaled features and stage-1 slopes of the CURRENT state.  They are carried across steps: the
  // stage-4 feature row of step h equals the stage-1 row of step h+1 except for the theta/gamma
  // columns 14, 15 (same P, V, A, theta_prev, gamma_prev), so x0 is x1 with two entries replaced,
  // and an equation that reads neither column has k1(h+1) == k4(h) bit for bit.
  T x0[18];
  cat_features18<T, M>(P0, P, V, A, th, ga, thp, gap, x0);
  cat_scale18<T>(x0);
  T k1t, k1g;
  cat_eval_pair<T, GENERIC, VM>(q, vm, x0, k1t, k1g);
  bool carry_t, carry_g;
  if constexpr (GENERIC || VM) {
    carry_t = q.carry_t != 0;
    carry_g = q.carry_g != 0;
  } else {
    carry_t = (CAT_DTHETA_LIVE_MASK & 0xC000u) == 0;
    carry_g = (CAT_DGAMMA_LIVE_MASK & 0xC000u) == 0;
  }

  // Controls are fetched ahead of their use and the reference point one step ahead, so the load latency hides behind
  // arithmetic.  One step of fp64 arithmetic (~3,000 cycles at 3.5 warps per scheduler) covers an HBM miss, so the
  // fp64 kernels load the next step's controls into registers.  One step of the fp32 kernel (~1,900 cycles, loads
  // issued ~40% into the step) does NOT (~1,000 cycles of DRAM latency at the natural clock; round-1 ncu: 3.2
  // long-scoreboard stall cycles per issued instruction): it keeps CAT_MPC_PF steps in flight through cp.async.
  S Vq[3] = {S(0), S(0), S(0)};
  int slot = 0;          // ASYNC: ring slot of step h
  int slot_in = 0;       // ASYNC: ring slot the next prefetch lands in
  CatArrive arrived = {-1, 0};
  if constexpr (ASYNC) {
    S* const mine = reinterpret_cast<S*>(s_ring.v) + threadIdx.x;
#pragma unroll
    for (int i = 0; i < CAT_MPC_PF; ++i) {
      if (i < q.H) {
        if constexpr (STREAM) cat_arrive_need<T>(q, arrived, i);
        S* d = mine + (i * 3) * CAT_MPC_BLOCK;
        cat_cp_async<S>(d, uc);
        cat_cp_async<S>(d + CAT_MPC_BLOCK, uc + q.sa);
        cat_cp_async<S>(d + 2 * CAT_MPC_BLOCK, uc + 2 * q.sa);
        uc += q.sh;
      }
      cat_cp_async_commit();   // one group per step, empty or not: the wait below counts groups
    }
    slot_in = CAT_MPC_PF;
  } else if constexpr (!RNG) {
    if constexpr (STREAM) {
      cat_arrive_need<T>(q, arrived, 0);
      Vq[0] = cat_ld_streamed<S>(uc);
      Vq[1] = cat_ld_streamed<S>(uc + q.sa);
      Vq[2] = cat_ld_streamed<S>(uc + 2 * q.sa);
    } else {
      Vq[0] = uc[0];
      Vq[1] = uc[q.sa];
      Vq[2] = uc[2 * q.sa];
    }
    uc += q.sh;
  }

#ifndef CAT_MPC_UNROLL
#define CAT_MPC_UNROLL 1   // A/B switch: 2 removes the ~38 loop-carried register moves per step, at twice the loop body
#endif
#define CAT_PRAGMA_(x) _Pragma(#x)
#define CAT_UNROLL_(n) CAT_PRAGMA_(unroll n)
  CAT_UNROLL_(CAT_MPC_UNROLL)
  for (int h = 0; h < q.H; ++h) {
    T Vn[3];
    if constexpr (RNG) {
      if (q.u_nom)
        cat_nominal_control<T, M>(q, seed_eff, prob, q.k_offset + kk, h, icdf, clip_r, Vn);
      else
        cat_sample_control<T, M>(seed_eff, (uint32_t)(q.problem_offset + prob), (uint64_t)(q.k_offset + kk),
                                 (uint32_t)h, q.sigma, clip_r, q.deep_clip != 0, icdf, Vn);
    } else if constexpr (ASYNC) {
      // the group of step h is the oldest of the CAT_MPC_PF in flight
      cat_cp_async_wait<CAT_MPC_PF - 1>();
      const S* mine = reinterpret_cast<const S*>(s_ring.v) + threadIdx.x + (slot * 3) * CAT_MPC_BLOCK;
      Vn[0] = T(mine[0]);
      Vn[1] = T(mine[CAT_MPC_BLOCK]);
      Vn[2] = T(mine[2 * CAT_MPC_BLOCK]);
      slot = slot == CAT_MPC_PF ? 0 : slot + 1;
    } else {
      Vn[0] = T(Vq[0]);
      Vn[1] = T(Vq[1]);
      Vn[2] = T(Vq[2]);
    }
    const T Rh[3] = {s_ref[3 * h], s_ref[3 * h + 1], s_ref[3 * h + 2]};
    if constexpr (ASYNC) {
      if (h + CAT_MPC_PF < q.H) {   // step h + PF into the slot that was read PF + 1 steps ago
        if constexpr (STREAM) cat_arrive_need<T>(q, arrived, h + CAT_MPC_PF);
        S* d = reinterpret_cast<S*>(s_ring.v) + threadIdx.x + (slot_in * 3) * CAT_MPC_BLOCK;
        cat_cp_async<S>(d, uc);
        cat_cp_async<S>(d + CAT_MPC_BLOCK, uc + q.sa);
        cat_cp_async<S>(d + 2 * CAT_MPC_BLOCK, uc + 2 * q.sa);
        uc += q.sh;
      }
      cat_cp_async_commit();
      slot_in = slot_in == CAT_MPC_PF ? 0 : slot_in + 1;
    } else if constexpr (!RNG) {
      if (h + 1 < q.H) {
        if constexpr (STREAM) {
          cat_arrive_need<T>(q, arrived, h + 1);
          Vq[0] = cat_ld_streamed<S>(uc);
          Vq[1] = cat_ld_streamed<S>(uc + q.sa);
          Vq[2] = cat_ld_streamed<S>(uc + 2 * q.sa);
        } else {
          Vq[0] = uc[0];
          Vq[1] = uc[q.sa];
          Vq[2] = uc[2 * q.sa];
        }
        uc += q.sh;
      }
    }
    T Pn[3], An[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      Pn[a] = P[a] + Vn[a] * q.dt_mm;
      An[a] = M::div_(Vn[a] - V[a], q.dt);
    }
    // ---- RK4 step of both angle ODEs (simulate_rk4_theta_gamma.py:61-66; k3 == k2) ----
    // rows: [0] = midpoint features (stages 2 and 3), [1] = end-point features (stage 4)
    T xs[2][18];
    cat_features18<T, M>(P0, Pn, Vn, An, th + q.dt * k1t, ga + q.dt * k1g, th, ga, xs[1]);
    cat_scale18<T>(xs[1]);
#pragma unroll
    for (int j = 0; j < 18; ++j) xs[0][j] = (x0[j] + xs[1][j]) * T(0.5);
    T kt[2], kg[2];
    cat_eval_pair_v<T, GENERIC, 2, VM>(q, vm, xs, kt, kg);
    const T k2t = kt[0], k4t = kt[1], k2g = kg[0], k4g = kg[1];
    // k1 + 2 k2 + 2 k3 + k4 with k3 == k2 (simulate_rk4_theta_gamma.py:65): 4 k2 is exact, one FMA
    const T thn = M::fma_(q.dt6, M::fma_(T(4), k2t, k1t + k4t), th);
    const T gan = M::fma_(q.dt6, M::fma_(T(4), k2g, k1g + k4g), ga);
    // ---- catenary terms of the new pose (main_fun.py:423) ----
    const T rx = Pn[0] - P0[0], ry = Pn[1] - P0[1], rz = Pn[2] - P0[2];
    const T l2 = rx * rx + ry * ry;
    const T l = M::sqrt_(l2);
    const T d2 = l2 + rz * rz;
    const CatSpan<T> sp = cat_span<T, M, false, true>(l, rz, q.L, q.c_lo, q.c_hi, warm);
    warm = sp.warm;
    // ---- cost, association as in the oracle ----
    const T ex = Pn[0] - Rh[0], ey = Pn[1] - Rh[1], ez = Pn[2] - Rh[2];
    const T dvx = Vn[0] - V[0], dvy = Vn[1] - V[1], dvz = Vn[2] - V[2];
    // The two penalty terms are evaluated only when one of them can be active (or an input is NaN) in some lane
    // of the warp -- a warp-uniform branch, so the inactive case issues nothing:
    //  * max(0, d - rho L) is zero iff d^2 <= (rho L)^2: no square root;
    //  * max(0, z_thr - z_low) is zero when a bound on the lowest point clears the threshold.  The cable from A
    //    over its lowest point Q to B is no shorter than the broken line A-Q-B, so (mirror B at Q's level)
    //    L^2 >= l^2 + (|dH| + 2t)^2 for Q at depth t below the lower end: z_low >= zmin - (sqrt(L^2-l^2) - |dH|)/2.
    //    With g = zmin - z_thr >= 0 the term vanishes once (2g + |dH|)^2 >= L^2 - l^2; no lowest point then.
    const T zmin = rz < T(0) ? P0[2] + rz : P0[2];
    const T g = zmin - q.z_thr;
    const T g2 = M::fma_(T(2), g, M::abs_(rz));
    const bool quiet = d2 <= q.rhoL2 && g >= T(0) && g2 * g2 >= q.L2 - l2;
    T over = T(0), deep = T(0);
    if (__any_sync(0xffffffffu, !quiet)) {
      over = M::pos_(M::sqrt_(d2) - q.rhoL);
      deep = M::pos_(q.z_thr - cat_z_low<T, M>(sp, l, q.L, P0[2], P0[2] + rz));
    }
    T c = q.w_trk * (ex * ex + ey * ey + ez * ez);
    c = c + q.w_du * (dvx * dvx + dvy * dvy + dvz * dvz);
    c = c + q.w_len * (over * over);
    c = c + q.w_sag * (deep * deep);
    c = c + q.w_ang * (thn * thn + gan * gan);
    if (!(sp.C == sp.C)) c = c + q.pen;
    J = J + c;
    thp = th;
    gap = ga;
    th = thn;
    ga = gan;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      P[a] = Pn[a];
      V[a] = Vn[a];
      A[a] = An[a];
    }
    // ---- stage-1 row and slopes of the next step ----
#pragma unroll
    for (int j = 0; j < 18; ++j) x0[j] = xs[1][j];
    x0[14] = cat_scale1<T>(th, 14);
    x0[15] = cat_scale1<T>(ga, 15);
    if constexpr (!GENERIC && !VM) {
      // shipped pair: which slopes carry over is known at compile time; only the other one is evaluated (an
      // unused sin still leaves its range check, branch and slow-path call behind -- the compiler cannot
      // prove the CUDA reduction routine free of side effects)
      constexpr bool CT = (CAT_DTHETA_LIVE_MASK & 0xC000u) == 0, CG = (CAT_DGAMMA_LIVE_MASK & 0xC000u) == 0;
      k1t = k4t;
      k1g = k4g;
      if constexpr (!CT) k1t = cat_dtheta_dt<T, M>(x0);
      if constexpr (!CG) k1g = cat_dgamma_dt<T, M>(x0);
    } else if (carry_t && carry_g) {
      k1t = k4t;
      k1g = k4g;
    } else {
      T nt, ng;
      cat_eval_pair<T, GENERIC, VM>(q, vm, x0, nt, ng);
      k1t = carry_t ? k4t : nt;
      k1g = carry_g ? k4g : ng;
    }
  }

  // Before a block touches anything the previous grid on the stream may still be using -- cost vector, per-block
  // winner slots, tickets, result buffers -- it waits for that grid to have completed (programmatic dependent
  // launch, see the top of the kernel); a whole horizon after its own start it has, so the wait is free.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (live && costs) costs[(long long)prob * q.K + k] = J;

  // ---- argmin over the candidates of this problem ----
  const double jc = live ? (double)J : __longlong_as_double(0x7ff0000000000000LL);
  const long long ji = live ? (q.k_offset + k) : 0x7fffffffffffffffLL;
  cat_grid_argmin(jc, ji, prob, block_best, tickets, best_cost, best_idx, q.p2p);
}

// fold of n winners, on the device: in = n x [cost bits, global idx] (the layout MpcDeviceStep.winner has and
// all_gather produces; also the chunk winners of the host pipeline), out = [cost bits, idx].  Input order, MPC
// winner order (cat_better_mpc) -> identical on every rank.
__device__ __forceinline__ void cat_fold_pairs(int n, const long long* __restrict__ in, long long& bc_bits,
                                               long long& bi) {
  double bc = __longlong_as_double(in[0]);
  bi = in[1];
  for (int g = 1; g < n; ++g) {
    const double c = __longlong_as_double(in[2 * g]);
    const long long i = in[2 * g + 1];
    if (cat_better_mpc(c, i, bc, bi)) {
      bc = c;
      bi = i;
    }
  }
  bc_bits = __double_as_longlong(bc);
}

static __global__ void k_reduce_winners(int n, const long long* __restrict__ in, long long* __restrict__ out) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  long long c, i;
  cat_fold_pairs(n, in, c, i);
  out[0] = c;
  out[1] = i;
}

// The same fold followed by the fused exchange (the host-pointer multi-GPU step with chunked explicit controls:
// the chunk launches carry no exchange, this one-block kernel does it).  <<<1, 64>>>
static __global__ void __launch_bounds__(64) k_p2p_exchange(int n, const long long* __restrict__ in, const CatP2PDev p2p,
                                                     double* __restrict__ best_cost,
                                                     long long* __restrict__ best_idx) {
  __shared__ long long s_mine[2];
  if (threadIdx.x == 0) cat_fold_pairs(n, in, s_mine[0], s_mine[1]);
  __syncthreads();
  cat_p2p_finish(s_mine[0], s_mine[1], p2p, best_cost, best_idx);
}

// job-wide winner of step p2p.seq without posting anything: closes a run of pipelined steps.  <<<1, 64>>>
static __global__ void __launch_bounds__(64) k_p2p_fold(const CatP2PDev p2p, double* __restrict__ best_cost,
                                                 long long* __restrict__ best_idx) {
  cat_p2p_fold(p2p, p2p.seq, best_cost, best_idx);
}

// materialise the on-device RNG controls: out[p][h][a][k] (SoA over candidates)
template <typename T>
__global__ void __launch_bounds__(256) k_sample_controls(const CatMpcDev<T> q, T* __restrict__ out) {
  typedef CatMath<T> M;
  __shared__ CatIcdfSmem<T> s_icdf;
  cat_icdf_stage<T>(s_icdf);
  const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= q.K) return;
  const int h = blockIdx.y, prob = blockIdx.z;
  T v[3];
  cat_sample_control<T, M>(q.seed, (uint32_t)(q.problem_offset + prob), (uint64_t)(q.k_offset + k), (uint32_t)h,
                           q.sigma, q.clip, q.deep_clip != 0, cat_icdf_ref_of<T>(s_icdf), v);
  T* o = out + (((long long)prob * q.H + h) * 3) * q.K + k;
  o[0] = v[0];
  o[q.K] = v[1];
  o[2 * q.K] = v[2];
}

// ------------------------------------------------------------------------------------------
// Peak microbenchmarks for the roofline denominators (MEASURED_PEAKS.json has HBM and bf16 only)
// kind 0: DFMA, 1: FFMA, 2: MUFU.EX2.  8 independent chains per thread.
// ------------------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(256) k_peak(int iters, double* sink) {
  if (KIND == 0) {
    double a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = 1.0 + 1e-9 * (threadIdx.x + j);
    const double m = 1.0000001, c = 1e-12;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = fma(a[j], m, c);
    }
    double s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += a[j];
    if (s == 12345.678) sink[0] = s;
  } else if (KIND == 1) {
    float a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = 1.0f + 1e-6f * (threadIdx.x + j);
    const float m = 1.0001f, c = 1e-6f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = fmaf(a[j], m, c);
    }
    float s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += a[j];
    if (s == 12345.678f) sink[0] = s;
  } else {
    float a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = 0.5f + 1e-3f * (threadIdx.x + j);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[j]));
    }
    float s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += a[j];
    if (s == 12345.678f) sink[0] = s;
  }
}
