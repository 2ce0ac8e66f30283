// Shared device/host helpers for libporl_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <vector>

#include "../../include/porl_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libporl_b200 is written for sm_100a (B200) only"
#endif

namespace porl {

constexpr int kMaxGroups = 8;       // problems per grouped contraction launch
constexpr int kSemInts = 1 << 16;   // split-K turnstile words
constexpr size_t kTcWsFloats = (size_t)1 << 24;   // 64 MB of split-K partial tiles for the tcgen05 kernel
constexpr int kPartialFloats = 1 << 20;  // block partials of the loss reductions
constexpr int kTicketBase = 64;          // counters[0] status, [1..63] kernel tickets, then per-chunk tickets
constexpr int kTickets = 4096;
constexpr int kGridBarBase = kTicketBase + kTickets;   // grid-barrier words of the multi-phase contraction launch
constexpr int kGridBarWords = 16;
constexpr int kCounterWords = kGridBarBase + kGridBarWords;

void set_error(const char* fmt, ...);

}  // namespace porl

struct porl_prof_rec {
  int cat;
  double work;
  cudaEvent_t e0, e1;
};

struct porl_ctx {
  bool profiling;
  std::vector<porl_prof_rec> prof;
  int device;
  int sm_count;
  int gemm_mode;            // 0 = fp32 FFMA, 1 = tcgen05 3xTF32 where supported
  int64_t launches;
  int64_t tc_launches;
  int* sem;                 // zero-initialised turnstile words (self-resetting)
  float* partials;          // loss reduction partials
  uint32_t* counters;       // [0] status word, [1..] last-block-done counters
  float* tc_ws;             // split-K partial tiles of the tcgen05 kernel (kTcWsFloats)
  int tc_a_in_tmem;         // tcgen05 kernel variant: A operand staged in tensor memory (default) or in shared memory
  int tc_chunk_kb;          // k-blocks per TMEM accumulation chunk of the tcgen05 kernel (1 or 2)
  int tc_phases;            // 1: consecutive contractions of an MLP pass share one persistent launch; 0 (default): one launch each
  int tc_terms;             // 3 = 3xTF32 (fp32-grade, product default); 1 = hi*hi only (plain TF32: the measured ceiling probe)
  unsigned long long* tc_dbg;  // optional per-CTA timestamps of the tcgen05 kernel (device memory), or null
  int pdl;                  // launch the step's kernels with programmatic stream serialization (default 1)
};

#define PORL_CHECK_ARG(cond, ...)                \
  do {                                           \
    if (!(cond)) {                               \
      porl::set_error(__VA_ARGS__);              \
      return PORL_ERR_INVALID;                   \
    }                                            \
  } while (0)

#define PORL_CUDA(call)                                                              \
  do {                                                                               \
    cudaError_t e_ = (call);                                                         \
    if (e_ != cudaSuccess) {                                                         \
      porl::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return PORL_ERR_CUDA;                                                          \
    }                                                                                \
  } while (0)

// Entry points launch on the CURRENT device with the caller's pointers: a context created for another device would
// silently run on the wrong GPU with foreign addresses (one process per GPU must cudaSetDevice before it builds
// anything).  Refuse instead.
#define PORL_ON_DEVICE(ctx)                                                          \
  do {                                                                               \
    int d_ = -1;                                                                     \
    if ((ctx) != nullptr && (cudaGetDevice(&d_) != cudaSuccess || d_ != (ctx)->device)) { \
      porl::set_error("%s: the context belongs to device %d but the current device is %d (call cudaSetDevice / " \
                      "torch.cuda.set_device first)", __func__, (ctx)->device, d_); \
      return PORL_ERR_INVALID;                                                       \
    }                                                                                \
  } while (0)

// after a kernel launch: count it and surface launch-configuration errors
#define PORL_LAUNCHED(ctx)                                                           \
  do {                                                                               \
    (ctx)->launches++;                                                               \
    cudaError_t e_ = cudaPeekAtLastError();                                          \
    if (e_ != cudaSuccess) {                                                         \
      porl::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e_), __FILE__, __LINE__); \
      return PORL_ERR_CUDA;                                                          \
    }                                                                                \
  } while (0)

// Brackets the launches issued during its lifetime with CUDA events when profiling is enabled.
struct ProfScope {
  porl_ctx* ctx;
  cudaStream_t st;
  size_t idx;
  bool on;
  ProfScope(porl_ctx* c, int cat, double work, cudaStream_t s) : ctx(c), st(s), idx(0), on(c->profiling) {
    if (!on) return;
    porl_prof_rec r;
    r.cat = cat;
    r.work = work;
    if (cudaEventCreate(&r.e0) != cudaSuccess || cudaEventCreate(&r.e1) != cudaSuccess) {
      on = false;
      return;
    }
    cudaEventRecord(r.e0, st);
    ctx->prof.push_back(r);
    idx = ctx->prof.size() - 1;
  }
  ~ProfScope() {
    if (on) cudaEventRecord(ctx->prof[idx].e1, st);
  }
};

static inline cudaStream_t as_stream(porl_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
static inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

#ifdef __CUDACC__
namespace porl {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Deterministic block sum for blockDim.x <= 1024 (multiple of 32); result valid in thread 0.
__device__ __forceinline__ float block_sum(float v, float* smem32) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) smem32[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float r = 0.f;
  if (warp == 0) {
    r = lane < nw ? smem32[lane] : 0.f;
    r = warp_sum(r);
  }
  return r;
}

// Programmatic dependent launch (PDL).  Every kernel launched through launch_pdl() starts with pdl_entry():
// it lets the NEXT kernel of the stream be scheduled as soon as all blocks of this grid are running (so its
// launch latency and prologue hide behind this grid's tail), then blocks until the PREVIOUS grid has completed and
// its writes are visible -- nothing of this kernel touches global memory before that, so the stream's order of
// effects is unchanged.  Because every such kernel waits, completion stays transitive along the chain.
__device__ __forceinline__ void pdl_entry() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// launch with the PDL attribute (ctx->pdl) -- the kernel MUST call pdl_entry() before its first global access
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(const porl_ctx* ctx, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = ctx->pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace porl
#endif
