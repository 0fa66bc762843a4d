// mpc_inst.cu -- instantiations of k_mpc_rollout for ONE (compute type, equation source) pair, selected by
//   -DCAT_INST_T=double|float  -DCAT_INST_KIND=0 (shipped pair, generated code) | 1 (any hall-of-fame row) |
//   2 (run-time equation programs).  build.py compiles this file six times, in parallel.
#include <cuda_runtime.h>

#include "cat_mpc_launch.cuh"

#if !defined(CAT_INST_T) || !defined(CAT_INST_KIND)
#error "compile with -DCAT_INST_T=<double|float> -DCAT_INST_KIND=<0|1|2>"
#endif

typedef CAT_INST_T InstT;
constexpr bool kGen = CAT_INST_KIND == CAT_KIND_GENERIC;
constexpr bool kVm = CAT_INST_KIND == CAT_KIND_VM;

template <>
int cat_mpc_launch_inst<InstT, CAT_INST_KIND>(const CatMpcLaunchArgs& a, const CatMpcDev<InstT>& q,
                                              const CatMpcProblem* problems, const void* controls,
                                              const double* p_ref, InstT* costs, CatBest* bb, unsigned int* tk,
                                              double* best_cost, long long* best_idx) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = a.grid;
  cfg.blockDim = dim3((unsigned)a.block, 1, 1);
  cfg.dynamicSmemBytes = a.smem;
  cfg.stream = a.stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = a.pdl ? 1 : 0;
  cudaError_t e;
  if (a.streamed) {
    // the streamed-upload form exists for the shipped equation pair only (the host step falls back to the chunked
    // upload for every other equation source)
#if CAT_INST_KIND == 0
    if (a.rng) return (int)cudaErrorInvalidValue;
    if (sizeof(InstT) == 8 && a.controls_f32)
      e = cudaLaunchKernelEx(&cfg, k_mpc_rollout<InstT, float, false, false, false, true>, q, problems,
                             (const float*)controls, p_ref, costs, bb, tk, best_cost, best_idx);
    else
      e = cudaLaunchKernelEx(&cfg, k_mpc_rollout<InstT, InstT, false, false, false, true>, q, problems,
                             (const InstT*)controls, p_ref, costs, bb, tk, best_cost, best_idx);
#else
    return (int)cudaErrorInvalidValue;
#endif
  } else if (a.rng) {
    e = cudaLaunchKernelEx(&cfg, k_mpc_rollout<InstT, InstT, kGen, kVm, true>, q, problems, (const InstT*)nullptr, p_ref,
                           costs, bb, tk, best_cost, best_idx);
  } else if (sizeof(InstT) == 8 && a.controls_f32) {
    e = cudaLaunchKernelEx(&cfg, k_mpc_rollout<InstT, float, kGen, kVm, false>, q, problems, (const float*)controls,
                           p_ref, costs, bb, tk, best_cost, best_idx);
  } else {
    e = cudaLaunchKernelEx(&cfg, k_mpc_rollout<InstT, InstT, kGen, kVm, false>, q, problems, (const InstT*)controls,
                           p_ref, costs, bb, tk, best_cost, best_idx);
  }
  if (e != cudaSuccess) return (int)e;
  return (int)cudaGetLastError();
}

template <> int cat_mpc_inst_stats<InstT, CAT_INST_KIND>(int enable, int reset, unsigned long long* out) {
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return (int)e;
  if (out) {
    unsigned long long v[4];
    e = cudaMemcpyFromSymbol(v, cat_dbg_counters, sizeof(v));
    if (e != cudaSuccess) return (int)e;
    for (int i = 0; i < 4; ++i) out[i] += v[i];
  }
  if (reset) {
    const unsigned long long z[4] = {0, 0, 0, 0};
    e = cudaMemcpyToSymbol(cat_dbg_counters, z, sizeof(z));
    if (e != cudaSuccess) return (int)e;
  }
  const int on = enable ? 1 : 0;
  e = cudaMemcpyToSymbol(cat_dbg_on, &on, sizeof(on));
  return (int)e;
}
