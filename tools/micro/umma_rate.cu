// Rate of tcgen05.mma kind::tf32 (M = 128, K = 8 per instruction) on one SM of a B200, in the conditions the 3xTF32
// contraction kernel (porl_b200/csrc/gemm_tc.cu) creates: 12 MMAs per 32-wide k-block into one TMEM accumulator,
//   * both operands from shared memory (SS) or A from tensor memory (TS),
//   * N = 128 or N = 256,
//   * alone, or with background warps that copy shared memory like the kernel's operand splitters (LDS.128, 12 ALU
//     ops, STS.128) and / or warps that drain a TMEM accumulator (tcgen05.ld) like its accumulator warps.
// Prints cycles per MMA and the shared-memory bytes per clock the background warps moved: the ceilings the main loop
// of the contraction kernel can be compared with.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/umma_rate tools/micro/umma_rate.cu
#include <cstdint>
#include <cstdio>
#include <string>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\tW_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra W_DONE;\n\tbra W_LOOP;\n\tW_DONE:\n\t}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// `leader` is the per-thread predicate (one lane of the converged warp issues); every other operand is warp-uniform,
// so the compiler keeps descriptors in uniform registers and emits one predicated UTCHMMA (no R2UR, no issue loop)
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc, uint32_t leader = 1) {
  asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
               "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
               "l"(a), "l"(b), "r"(idesc), "r"(acc), "r"(leader)
               : "memory");
}
__device__ __forceinline__ void mma_ss_f16(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc, uint32_t leader = 1) {
  asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
               "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
               "l"(a), "l"(b), "r"(idesc), "r"(acc), "r"(leader)
               : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc, uint32_t leader = 1) {
  asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
               "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
               "r"(a), "l"(b), "r"(idesc), "r"(acc), "r"(leader)
               : "memory");
}
__device__ __forceinline__ void tc_commit_pred(uint32_t bar, uint32_t leader) {
  asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %1, 0;\n\t"
               "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar), "r"(leader)
               : "memory");
}
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr) {   // K-major SWIZZLE_128B, SBO 1024
  return (uint64_t)((addr >> 4) & 0x3fffu) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xffffffff;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
  return pred != 0;
}
__device__ __forceinline__ float split_lo(float x) {
  const float d = x - __uint_as_float(__float_as_uint(x) & 0xffffe000u);
  return __uint_as_float(__float_as_uint(d) + 0x1000u);
}

struct Cfg {
  int n;          // MMA N: 128 or 256
  int a_tmem;     // 0: SS, 1: TS
  int n_acc;      // accumulators used round-robin per k-block (1 or 2; N = 256: 1 when a_tmem)
  int mmas;       // MMAs per k-block (12 = 3xTF32, 4 = one pass)
  int bg_warps;   // 0..6 background copy warps (warps 2..7)
  int drain;      // 1: warps 8..11 drain 128 TMEM columns per loop, back to back
  int inflight;   // k-blocks the issuer keeps in flight (commit ring)
  int iters;      // k-blocks
  int f16;        // 1: kind::f16 (bf16 operands, K = 16 per instruction) instead of kind::tf32
  int alt;        // 1: consecutive MMAs alternate between two accumulators (no back-to-back dependency)
  int conv;       // 1: the whole warp runs the issue loop converged, one lane predicated (uniform-register descriptors)
  int restart;    // 1: every k-block starts a fresh accumulation (first MMA does not accumulate); 0: only the first k-block
  int commit_every;   // commit after every n-th k-block (1 = the kernel's pattern)
};

constexpr int kStage = 65536, kStages = 3;
constexpr int kSmem = 1024 + kStages * kStage + 256;

__global__ void __launch_bounds__(384, 1) umma_kernel(Cfg c, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw), base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  uint64_t* bars = reinterpret_cast<uint64_t*>(base_ptr + kStages * kStage);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  volatile int* done = reinterpret_cast<volatile int*>(tmem_slot + 1);
  const uint32_t bar0 = smem_u32(bars);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < kStages * kStage / 4; i += blockDim.x)
    reinterpret_cast<float*>(base_ptr)[i] = 1.0f + 1e-3f * (float)(i & 1023);
  if (threadIdx.x == 0) {
    for (int b = 0; b < 8; ++b) mbar_init(bar0 + 8u * b, 1);
    *done = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  const int warp_u = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);     // provably warp-uniform
  if (warp_u == 1 && c.conv) {
    // ---- converged issuer: all 32 lanes run the loop with warp-uniform operands, one elected lane issues
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t*>(tmem_slot), 0);
    const bool leader = elect_one();
    const uint32_t fmt = c.f16 ? 1u : 2u;
    const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(c.n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const long long t0 = clock64();
    const int groups = c.mmas / 4;
    for (int kb = 0; kb < c.iters; ++kb) {
      if (c.commit_every == 1 && kb >= c.inflight)
        mbar_wait(bar0 + 8u * ((kb - c.inflight) & 7), (uint32_t)(((kb - c.inflight) >> 3) & 1));
      const uint32_t st = base + (uint32_t)(kb % kStages) * kStage;
      const uint64_t a_hi = smem_desc(st), b_hi = smem_desc(st + 16384u);
      const uint32_t d = tmem_base + (uint32_t)((kb % c.n_acc) * c.n);
      const uint32_t at = tmem_base + 256u + (uint32_t)((kb & 3) * 64);
      uint32_t acc = (c.restart || kb == 0) ? 0u : 1u;
      for (int gidx = 0; gidx < groups; ++gidx) {
        const int plane = gidx % 3;
        const uint64_t a0 = a_hi + (plane == 0 ? (32768u >> 4) : 0u);
        const uint64_t b0 = b_hi + (plane == 1 ? ((c.n == 256 ? 16384u : 32768u) >> 4) : 0u);
        const uint32_t at0 = at + (plane == 0 ? 32u : 0u);
        if (leader) {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            if (c.f16) mma_ss_f16(d, a0 + 2u * ks, b0 + 2u * ks, idesc, ks ? 1u : acc);
            else if (c.a_tmem) mma_ts(d, at0 + 8u * ks, b0 + 2u * ks, idesc, ks ? 1u : acc);
            else mma_ss(d, a0 + 2u * ks, b0 + 2u * ks, idesc, ks ? 1u : acc);
          }
        }
        acc = 1;
      }
      if (leader && ((kb + 1) % c.commit_every == 0 || kb == c.iters - 1)) tc_commit(bar0 + 8u * (kb & 7));
      __syncwarp();
    }
    if (c.commit_every == 1) {
      for (int kb = (c.iters > c.inflight ? c.iters - c.inflight : 0); kb < c.iters; ++kb)
        mbar_wait(bar0 + 8u * (kb & 7), (uint32_t)((kb >> 3) & 1));
    } else {                    // phases are not tracked in this mode: spin on the clock long enough for the tail to drain
      const long long tw = clock64();
      while (clock64() - tw < 20000) {}
    }
    const long long t1 = clock64() - (c.commit_every == 1 ? 0 : 20000);
    if (lane == 0) {
      out[blockIdx.x * 4 + 0] = t1 - t0;
      *done = 1;
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // D fp32; A/B format 2 = tf32 (kind::tf32) or 1 = bf16 (kind::f16)
      const uint32_t fmt = c.f16 ? 1u : 2u;
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(c.n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const long long t0 = clock64();
      for (int kb = 0; kb < c.iters; ++kb) {
        if (kb >= c.inflight) mbar_wait(bar0 + 8u * ((kb - c.inflight) & 7), (uint32_t)(((kb - c.inflight) >> 3) & 1));
        const uint32_t st = base + (uint32_t)(kb % kStages) * kStage;
        const uint64_t a_hi = smem_desc(st), b_hi = smem_desc(st + 16384u);
        const uint32_t d = tmem_base + (uint32_t)((kb % c.n_acc) * c.n);
        const uint32_t at = tmem_base + 256u + (uint32_t)((kb & 3) * 64);
        uint32_t acc = 0, acc2 = 0;
        for (int j = 0; j < c.mmas; ++j) {
          const int ks = j & 3, plane = (j >> 2) % 3;                 // plane 0: lo*hi, 1: hi*lo, 2: hi*hi (distinct smem tiles)
          const uint64_t a = a_hi + 2u * ks + (plane == 0 ? (32768u >> 4) : 0u);
          const uint64_t b = b_hi + 2u * ks + (plane == 1 ? ((c.n == 256 ? 16384u : 32768u) >> 4) : 0u);
          if (c.alt && (j & 1)) {
            mma_ss(d ^ 128u, a, b, idesc, acc2);
            acc2 = 1;
            continue;
          }
          if (c.f16) mma_ss_f16(d, a, b, idesc, acc);
          else if (c.a_tmem) mma_ts(d, at + 8u * ks + (plane == 0 ? 32u : 0u), b, idesc, acc);
          else mma_ss(d, a, b, idesc, acc);
          acc = 1;
        }
        tc_commit(bar0 + 8u * (kb & 7));
        if (c.mmas == 0) {                     // background-only run: hold each k-block for 1000 clocks
          const long long tw = clock64();
          while (clock64() - tw < 1000) {}
        }
      }
      for (int kb = (c.iters > c.inflight ? c.iters - c.inflight : 0); kb < c.iters; ++kb)
        mbar_wait(bar0 + 8u * (kb & 7), (uint32_t)((kb >> 3) & 1));
      const long long t1 = clock64();
      out[blockIdx.x * 4 + 0] = t1 - t0;
      *done = 1;
    }
  } else if (warp >= 2 && warp < 8) {
    long long moved = 0;
    if (warp - 2 < c.bg_warps) {
      const int t = threadIdx.x - 64, nthr = c.bg_warps * 32;
      int s = 0;
      while (!*done) {
        const float4* src = reinterpret_cast<const float4*>(base_ptr + s * kStage);
        float4* dst = reinterpret_cast<float4*>(base_ptr + s * kStage + 32768 + 16384 + 8192);   // scratch part of the stage
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int e = (t + i * nthr) & 511;                     // 8 KB window
          const float4 v = src[e + 512 * (i & 3)];
          dst[e] = make_float4(split_lo(v.x), split_lo(v.y), split_lo(v.z), split_lo(v.w));
        }
        moved += 8 * 32;            // bytes read + written per thread
        s = (s + 1) % kStages;
      }
    }
    if (lane == 0 && warp == 2) out[blockIdx.x * 4 + 1] = moved * 32 * (long long)c.bg_warps;
  } else if (warp >= 8) {
    long long loops = 0;
    if (c.drain) {
      float accv[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) accv[i] = 0.f;
      const uint32_t tb = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + 128u;      // second accumulator's columns
      while (!*done) {
#pragma unroll
        for (int sl = 0; sl < 4; ++sl) {
          float v[32];
          tc_ld32(tb + 32u * sl, v);
#pragma unroll
          for (int i = 0; i < 32; ++i) accv[i] += v[i];
        }
        ++loops;
      }
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 32; ++i) s += accv[i];
      if (s == 123.456f) out[blockIdx.x * 4 + 3] = 1;
    }
    if (lane == 0 && warp == 8) out[blockIdx.x * 4 + 2] = loops;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
}

// The contraction kernel's issue pattern with nothing else in the loop: compile-time shape, 12 MMAs (lo*hi, hi*lo, hi*hi
// over four 8-wide k-steps) + one commit per k-block, at most two k-blocks in flight, converged issuing warp.  This is
// what --json reports: the rate of the MMA form itself.
template <int N, bool TS>
__global__ void __launch_bounds__(128, 1) umma_lean_kernel(int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw), base = (raw + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - raw);
  uint64_t* bars = reinterpret_cast<uint64_t*>(base_ptr + kStages * kStage);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  const uint32_t bar0 = smem_u32(bars);
  for (int i = threadIdx.x; i < kStages * kStage / 4; i += blockDim.x) reinterpret_cast<float*>(base_ptr)[i] = 1.0f;
  if (threadIdx.x == 0) {
    for (int b = 0; b < 8; ++b) mbar_init(bar0 + 8u * b, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  const int warp_u = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  if (warp_u == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (warp_u == 1) {
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t*>(tmem_slot), 0);
    const bool leader = elect_one();
    constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    constexpr uint32_t b_lo_off = (N == 256 ? 16384u : 32768u) >> 4;
    const long long t0 = clock64();
    for (int kb = 0; kb < iters; ++kb) {
      if (kb >= 2) mbar_wait(bar0 + 8u * ((kb - 2) & 7), (uint32_t)(((kb - 2) >> 3) & 1));
      const uint32_t st = base + (uint32_t)(kb % kStages) * kStage;
      const uint64_t a_hi = smem_desc(st), b_hi = smem_desc(st + 16384u);
      const uint32_t d = tmem_base + (uint32_t)((kb & 1) * (N == 256 ? 0 : 128));
      const uint32_t at = tmem_base + 256u + (uint32_t)((kb & 3) * 64);
      if (leader) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          if (TS) mma_ts(d, at + 32u + 8u * ks, b_hi + 2u * ks, idesc, ks ? 1u : 0u);
          else mma_ss(d, a_hi + (32768u >> 4) + 2u * ks, b_hi + 2u * ks, idesc, ks ? 1u : 0u);
          if (TS) mma_ts(d, at + 8u * ks, b_hi + b_lo_off + 2u * ks, idesc, 1u);
          else mma_ss(d, a_hi + 2u * ks, b_hi + b_lo_off + 2u * ks, idesc, 1u);
        }
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          if (TS) mma_ts(d, at + 8u * ks, b_hi + 2u * ks, idesc, 1u);
          else mma_ss(d, a_hi + 2u * ks, b_hi + 2u * ks, idesc, 1u);
        }
        tc_commit(bar0 + 8u * (kb & 7));
      }
      __syncwarp();
    }
    for (int kb = (iters > 2 ? iters - 2 : 0); kb < iters; ++kb) mbar_wait(bar0 + 8u * (kb & 7), (uint32_t)((kb >> 3) & 1));
    const long long t1 = clock64();
    if (threadIdx.x == 32) out[blockIdx.x * 4] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  if (warp_u == 1) {
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

template <int N, bool TS>
static double lean_clk_per_mma(long long* out) {
  const int iters = 1000;
  cudaFuncSetAttribute(umma_lean_kernel<N, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  for (int rep = 0; rep < 2; ++rep) umma_lean_kernel<N, TS><<<148, 128, kSmem>>>(iters, out);
  if (cudaDeviceSynchronize() != cudaSuccess) return -1.0;
  long long h[148 * 4];
  cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
  double clk = 0;
  for (int i = 0; i < 148; ++i) clk += (double)h[4 * i] / 148;
  return clk / iters / 12.0;
}

// --json: the four forms the bench quotes (12 MMAs + one commit per k-block, converged issuer), as one JSON line
static int quick_json(long long* out) {
  const double v[4] = {lean_clk_per_mma<128, false>(out), lean_clk_per_mma<256, false>(out), lean_clk_per_mma<128, true>(out),
                       lean_clk_per_mma<256, true>(out)};
  const char* names[] = {"ss_n128", "ss_n256", "ts_n128", "ts_n256"};
  const int n[4] = {128, 256, 128, 256};
  printf("{");
  for (int k = 0; k < 4; ++k) {
    if (v[k] <= 0) return 1;
    printf("%s\"%s_clk_per_mma\": %.1f, \"%s_fma_per_clk_sm\": %.0f", k ? ", " : "", names[k], v[k], names[k], 128.0 * n[k] * 8 / v[k]);
  }
  printf("}\n");
  return 0;
}

int main(int argc, char** argv) {
  long long* out;
  cudaMalloc(&out, 148 * 4 * sizeof(long long));
  cudaFuncSetAttribute(umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  if (argc > 1 && std::string(argv[1]) == "--json") return quick_json(out);
  const Cfg cfgs[] = {
      // n, a_tmem, n_acc, mmas, bg, drain, inflight, iters, f16, alt, conv, restart, commit_every
      {128, 0, 1, 12, 0, 0, 2, 2000, 0, 0, 0, 1, 1},  {128, 0, 1, 12, 0, 0, 2, 2000, 0, 0, 1, 1, 1},  {128, 0, 1, 4, 0, 0, 2, 2000, 0, 0, 1, 1, 1},
      {128, 0, 1, 96, 0, 0, 2, 250, 0, 0, 1, 1, 1},   {128, 0, 1, 12, 0, 0, 2, 2000, 0, 0, 1, 0, 1},  {128, 0, 1, 12, 0, 0, 2, 2000, 0, 0, 1, 1, 8},
      {256, 0, 1, 12, 0, 0, 2, 2000, 0, 0, 1, 1, 1},  {256, 0, 1, 96, 0, 0, 2, 250, 0, 0, 1, 1, 1},   {64, 0, 1, 12, 0, 0, 2, 2000, 0, 0, 1, 1, 1},
      {128, 1, 2, 12, 0, 0, 2, 2000, 0, 0, 1, 1, 1},  {256, 1, 1, 12, 0, 0, 2, 2000, 0, 0, 1, 1, 1},  {128, 0, 1, 12, 0, 0, 2, 2000, 1, 0, 1, 1, 1},
      {256, 0, 1, 12, 0, 0, 2, 2000, 1, 0, 1, 1, 1},  {256, 0, 1, 96, 0, 0, 2, 250, 1, 0, 1, 1, 1},
      {128, 0, 2, 12, 2, 0, 2, 2000, 0, 0, 1, 1, 1},  {128, 0, 2, 12, 4, 0, 2, 2000, 0, 0, 1, 1, 1},  {128, 0, 2, 12, 6, 0, 2, 2000, 0, 0, 1, 1, 1},
      {128, 1, 2, 12, 6, 0, 2, 2000, 0, 0, 1, 1, 1},  {256, 0, 1, 12, 6, 0, 2, 2000, 0, 0, 1, 1, 1},  {256, 1, 1, 12, 6, 0, 2, 2000, 0, 0, 1, 1, 1},
      {128, 0, 1, 12, 0, 1, 2, 2000, 0, 0, 1, 1, 1},  {128, 0, 1, 12, 6, 1, 2, 2000, 0, 0, 1, 1, 1},  {128, 1, 1, 12, 6, 1, 2, 2000, 0, 0, 1, 1, 1},
      {128, 0, 2, 0, 6, 0, 2, 2000, 0, 0, 0, 1, 1},
  };




  printf("tcgen05.mma kind::tf32 M=128 K=8, one issuing thread per SM, 148 CTAs; k-block = `mmas` MMAs then a commit\n");
  printf("%-4s %-3s %-4s %-4s %-5s %-5s %-3s %-6s %-9s | %12s %12s %14s %14s %16s\n", "N", "A", "kind", "conv", "n_acc", "mmas", "bg", "drain", "inflight",
         "clk/kblock", "clk/MMA", "MMA FMA/clk/SM", "bg smem B/clk", "drain KB/kblock");
  for (const Cfg& c : cfgs) {
    for (int rep = 0; rep < 2; ++rep) umma_kernel<<<148, 384, kSmem>>>(c, out);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148 * 4];
    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    double clk = 0, moved = 0, loops = 0;
    for (int i = 0; i < 148; ++i) {
      clk += (double)h[4 * i] / 148;
      moved += (double)h[4 * i + 1] / 148;
      loops += (double)h[4 * i + 2] / 148;
    }
    const double per_kb = clk / c.iters, per_mma = c.mmas ? per_kb / c.mmas : 0.0;
    printf("%-4d %-3s %-4s %-4d %-5d %-5d %-3d %-6d %-9d | %12.1f %12.1f %14.0f %14.1f %16.1f  [%s]\n", c.n, c.a_tmem ? "TS" : "SS", c.f16 ? "bf16" : "tf32", c.conv, c.n_acc,
           c.mmas, c.bg_warps, c.drain, c.inflight, per_kb, per_mma, per_mma > 0 ? 128.0 * c.n * (c.f16 ? 16 : 8) / per_mma : 0.0, moved / clk,
           loops * 64.0 / c.iters, cudaGetErrorString(e));
  }
  return 0;
}
