// cat_host.cuh -- host-side plumbing shared by the translation units of libcatenary_b200.so:
// the context (device, streams, grow-only scratch), error reporting, launch bookkeeping, NVTX ranges.
// Everything here has internal linkage; each translation unit carries its own copy of the helpers.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/catenary_b200.h"

enum { SLOT_A = 0, SLOT_B, SLOT_C, SLOT_D, SLOT_E, SLOT_F, SLOT_G, SLOT_H, SLOT_BEST, SLOT_TICKET, SLOT_ARRIVE, N_SLOTS };

struct CatHostReg {   // one cudaHostRegister'ed caller range (cat_host_register)
  void* p;
  size_t bytes;
};

struct CatCtx {
  int device;
  cudaStream_t own_stream;
  cudaStream_t aux_stream;    // second lane of the chunked pipelines
  cudaStream_t stream;        // stream in use (own_stream or a caller-owned one)
  cudaEvent_t ev0, ev1, ev_aux;
  cudaEvent_t ev_chunk[8];    // H2D chunk k of the pipelined host MPC step is on the device
  std::string err;
  int64_t launches;
  void* slot[N_SLOTS];
  size_t cap[N_SLOTS];
  int sm_count;
  int mpc_block;   // 0 = heuristic; else forced block size (CAT_MPC_BLOCK_SIZE env, tuning only)
  // device-mapped pinned staging of the host MPC step: [0, kStageWin) winners written by the kernel's last block
  // (no D2H copy), [kStageWin, kStageBytes) problems + reference rows going up in ONE async copy
  unsigned char* h_stage;
  unsigned char* d_stage;   // device alias of h_stage
  void* ws;                 // CatMpcWs* selected by cat_set_workspace; null = the shared scratch slots
  int overlap;              // rollouts launch with programmatic stream serialization (cat_set_overlap; default on)
  int nvtx;                 // CAT_NVTX=1: NVTX ranges around step / exchange / host copies
  CatHostReg reg[64];
  int n_reg;
  int streamed;               // host MPC step: candidates stream in under the running rollout (CAT_MPC_STREAMED=0: off)
  unsigned int arrive_epoch;  // value the streamed upload's arrival words take in the current call
  void* write_value32;        // cuStreamWriteValue32 (driver entry point, resolved in cat_init) or null
  int chunk_n;                // chunked upload: equal chunks ahead of the last one (CAT_MPC_CHUNK_N, tuning only)
  long long chunk_last;       // ... candidates in the last chunk, 0 = one block per SM (CAT_MPC_CHUNK_LAST)
  int stream_split[8];        // CAT_MPC_STREAM_SPLIT=a,b,c: row-group sizes in horizon steps (tuning only); [0] = 0: rule
};
constexpr size_t kStageWin = 4096, kStageBytes = 65536;

namespace {

inline int fail(CatCtx* c, int code, const char* what, cudaError_t e = cudaSuccess) {
  if (c) {
    c->err = what;
    if (e != cudaSuccess) {
      c->err += ": ";
      c->err += cudaGetErrorString(e);
    }
  }
  return code;
}

#define CU(call)                                              \
  do {                                                        \
    cudaError_t e__ = (call);                                 \
    if (e__ != cudaSuccess) return fail(c, CAT_ECUDA, #call, e__); \
  } while (0)

#define NEED(cond)                                            \
  do {                                                        \
    if (!(cond)) return fail(c, CAT_EINVAL, "invalid argument: " #cond); \
  } while (0)

inline int bind(CatCtx* c) {
  CU(cudaSetDevice(c->device));
  return CAT_OK;
}

// grow-only device scratch; zero-filled when (re)allocated if zero == true.  Regrowing frees the old buffer,
// which work enqueued earlier on ANY stream of this device may still be using (a caller can move the context
// between streams with cat_set_stream), so the whole device is drained first -- it happens a handful of times
// in the life of a context.
inline int scratch(CatCtx* c, int s, size_t bytes, void** out, bool zero = false) {
  if (bytes == 0) bytes = 16;
  if (c->cap[s] < bytes) {
    if (c->slot[s]) {
      CU(cudaDeviceSynchronize());
      CU(cudaFree(c->slot[s]));
      c->slot[s] = nullptr;
      c->cap[s] = 0;
    }
    size_t want = bytes + (bytes >> 2);
    cudaError_t e = cudaMalloc(&c->slot[s], want);
    if (e != cudaSuccess) {
      want = bytes;
      e = cudaMalloc(&c->slot[s], want);
    }
    if (e != cudaSuccess) {
      c->slot[s] = nullptr;
      return fail(c, CAT_ENOMEM, "cudaMalloc", e);
    }
    c->cap[s] = want;
    if (zero) {
      CU(cudaMemsetAsync(c->slot[s], 0, want, c->stream));
      CU(cudaStreamSynchronize(c->stream));   // visible to every stream the context may be moved to
    }
  }
  *out = c->slot[s];
  return CAT_OK;
}

#define SCRATCH(slot_, bytes_, ptr_)                                   \
  do {                                                                 \
    void* p__ = nullptr;                                               \
    int rc__ = scratch(c, slot_, (size_t)(bytes_), &p__);              \
    if (rc__ != CAT_OK) return rc__;                                   \
    ptr_ = reinterpret_cast<decltype(ptr_)>(p__);                      \
  } while (0)

#define LAUNCHED()                                                     \
  do {                                                                 \
    ++c->launches;                                                     \
    cudaError_t e__ = cudaGetLastError();                              \
    if (e__ != cudaSuccess) return fail(c, CAT_ECUDA, "kernel launch", e__); \
  } while (0)

inline unsigned int nblk(long long n, int b) { return (unsigned int)((n + b - 1) / b); }

}  // namespace

// NVTX ranges (SURVEY.md section 5 "tracing"): off unless CAT_NVTX=1 -- the header-only nvtx3 resolves the
// injection library at run time, so nothing is linked and an un-profiled run pays one predictable branch.
void cat_nvtx_push(const CatCtx* c, const char* name);
void cat_nvtx_pop(const CatCtx* c);
struct CatNvtxRange {
  const CatCtx* c;
  CatNvtxRange(const CatCtx* ctx, const char* name) : c(ctx) { if (c && c->nvtx) cat_nvtx_push(c, name); }
  ~CatNvtxRange() { if (c && c->nvtx) cat_nvtx_pop(c); }
};
