// cat_mpc_launch.cuh -- seam between the C ABI (catenary_b200.cu) and the translation units that instantiate
// k_mpc_rollout (mpc_inst.cu, compiled once per (compute type, equation source) so the six heavy instantiation sets
// build in parallel).
#pragma once
#include "cat_kernels_mpc.cuh"

enum { CAT_KIND_SHIPPED = 0, CAT_KIND_GENERIC = 1, CAT_KIND_VM = 2 };

struct CatMpcLaunchArgs {
  dim3 grid;
  int block;
  size_t smem;
  cudaStream_t stream;
  bool rng;          // controls drawn on the device
  bool controls_f32; // explicit controls stored as float while T = double
  bool pdl;          // programmatic stream serialization: may start under the tail of the preceding rollout
  bool streamed;     // controls arrive while the kernel runs (q.arrive; shipped equations, explicit controls only)
};

// one launch of k_mpc_rollout<T, S, ...>; returns the cudaError_t of the launch
template <typename T, int KIND>
int cat_mpc_launch_inst(const CatMpcLaunchArgs& a, const CatMpcDev<T>& q, const CatMpcProblem* problems,
                        const void* controls, const double* p_ref, T* costs, CatBest* bb, unsigned int* tk,
                        double* best_cost, long long* best_idx);
// cold-path counters of that translation unit (cat_math.cuh cat_dbg_*): set the switch, add the counters to
// out[4], optionally zero them.  Returns a cudaError_t.
template <typename T, int KIND> int cat_mpc_inst_stats(int enable, int reset, unsigned long long* out);

#define CAT_MPC_INST_DECL(T, KIND)                                                                                 \
  template <>                                                                                                       \
  int cat_mpc_launch_inst<T, KIND>(const CatMpcLaunchArgs&, const CatMpcDev<T>&, const CatMpcProblem*, const void*, \
                                   const double*, T*, CatBest*, unsigned int*, double*, long long*);               \
  template <> int cat_mpc_inst_stats<T, KIND>(int, int, unsigned long long*);
CAT_MPC_INST_DECL(double, 0)
CAT_MPC_INST_DECL(double, 1)
CAT_MPC_INST_DECL(double, 2)
CAT_MPC_INST_DECL(float, 0)
CAT_MPC_INST_DECL(float, 1)
CAT_MPC_INST_DECL(float, 2)
#undef CAT_MPC_INST_DECL
