// FP64 pipe microbenchmark for sm_100a: dependent-issue latency of DFMA / DMUL / DADD / DSETP+select and the
// throughput reached by W warps per SM sub-partition with ILP independent chains each.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_latency fp64_latency.cu && ./fp64_latency
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void k_dfma(double* out, long long* cyc, int n, double a, double b) {
  double x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = threadIdx.x * 1e-3 + i;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < n; ++it) {
#pragma unroll
    for (int r = 0; r < 16; ++r)
#pragma unroll
      for (int i = 0; i < ILP; ++i) x[i] = fma(x[i], a, b);
  }
  const long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// one DFMA of a dependent chain + NI independent ops of another kind per step (volatile asm: not collapsible)
// KIND 0 IMAD (mad.lo.s32)  1 FFMA  2 IADD3 (add.s32)  3 LOP3 (xor.b32)  4 SEL (selp)  5 MUFU (ex2.approx)
// 6 shared-memory load  7 uniform-datapath/constant load (ld.const)
template <int NI, int KIND>
__global__ void k_mixed(double* out, long long* cyc, int n, double a, double b) {
  __shared__ int sm[32];
  double x = threadIdx.x * 1e-3;
  int j[4] = {(int)threadIdx.x, 1, 2, 3};
  float f[4] = {1.f, 2.f, 3.f, 4.f};
  if (threadIdx.x < 32) sm[threadIdx.x] = threadIdx.x;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < n; ++it) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      x = fma(x, a, b);
#pragma unroll
      for (int i = 0; i < NI; ++i) {
        int& J = j[i & 3];
        float& F = f[i & 3];
        if (KIND == 0) asm volatile("mad.lo.s32 %0, %0, 3, 1;" : "+r"(J));
        if (KIND == 1) asm volatile("fma.rn.f32 %0, %0, 0f3F800347, 0f3F000000;" : "+f"(F));
        if (KIND == 2) asm volatile("add.s32 %0, %0, 7;" : "+r"(J));
        if (KIND == 3) asm volatile("xor.b32 %0, %0, 0x5a5a;" : "+r"(J));
        if (KIND == 4) asm volatile("{.reg .pred p; setp.gt.s32 p, %0, 5; selp.b32 %0, %0, 9, p;}" : "+r"(J));
        if (KIND == 5) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(F));
        if (KIND == 6) asm volatile("ld.shared.b32 %0, [%1];" : "=r"(J) : "r"((unsigned)__cvta_generic_to_shared(&sm[(J + i) & 31])));
      }
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x + j[0] + j[1] + j[2] + j[3] + f[0] + f[1] + f[2] + f[3];
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
  double* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 20);
  cudaMalloc(&cyc, 1 << 12);
  const int n = 2048;
  long long h;
  printf("warps/SMSP  ILP  cycles per DFMA per warp   DFMA per cycle per SMSP (peak 0.5)\n");
  for (int w : {1, 2, 3, 4, 6, 8}) {
    for (int ilp : {1, 2, 4}) {
      for (int rep = 0; rep < 2; ++rep) {
        if (ilp == 1) k_dfma<1><<<1, 128 * w>>>(out, cyc, n, 1.0000001, 1e-9);
        if (ilp == 2) k_dfma<2><<<1, 128 * w>>>(out, cyc, n, 1.0000001, 1e-9);
        if (ilp == 4) k_dfma<4><<<1, 128 * w>>>(out, cyc, n, 1.0000001, 1e-9);
        cudaDeviceSynchronize();
      }
      cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      const double per = (double)h / (n * 16.0 * ilp);
      printf("%6d %6d %12.2f %24.3f\n", w, ilp, per, w / per);
    }
  }
  printf("4 warps/SMSP, each: 1 chained DFMA + NI independent ops per step -> cycles per step per warp (8 = FP64 pipe bound)\n");
#define RUN(NI, KIND)                                                      \
  for (int rep = 0; rep < 2; ++rep) {                                      \
    k_mixed<NI, KIND><<<1, 512>>>(out, cyc, n, 1.0000001, 1e-9);           \
    cudaDeviceSynchronize();                                               \
  }                                                                        \
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);                          \
  printf("  kind %d NI %d: %.2f\n", KIND, NI, (double)h / (n * 16.0));
  RUN(0, 0)
  RUN(1, 0) RUN(2, 0) RUN(4, 0)
  RUN(1, 1) RUN(2, 1) RUN(4, 1)
  RUN(1, 2) RUN(2, 2) RUN(4, 2)
  RUN(1, 3) RUN(2, 3) RUN(4, 3)
  RUN(1, 4) RUN(2, 4) RUN(4, 4)
  RUN(1, 5) RUN(2, 5) RUN(4, 5)
  RUN(1, 6) RUN(2, 6) RUN(4, 6)
  printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
