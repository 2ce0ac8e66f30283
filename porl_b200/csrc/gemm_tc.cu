// tcgen05 3xTF32 grouped contraction: fp32-grade accuracy on the 5th-gen tensor cores.
//
// Every operand arrives as a (hi, lo) pair of tf32-exact planes (gemm.cuh).  One output tile
// C[128 x 128] per CTA; per 32-wide k-block TMA stages four swizzled tiles (A_hi, A_lo, B_hi, B_lo,
// 16 KB each) and one elected thread issues, per 8-wide k-step, three tcgen05.mma.kind::tf32 into a
// single fp32 TMEM accumulator:   A_lo*B_hi + A_hi*B_lo + A_hi*B_hi   (the A_lo*B_lo term, 2^-22
// relative, is dropped).  Warp roles: 0 = TMA producer, 1 = TMEM allocator + MMA issuer,
// 2..5 = epilogue (tcgen05.ld -> smem transpose -> coalesced global stores with bias / ReLU / tanh /
// ReLU-mask / pair split fused).  mbarrier ring: full[s] (TMA -> MMA), empty[s] (tcgen05.commit -> TMA),
// tmem_full (last commit -> epilogue).  Split-K uses the same turnstile words as the FFMA kernel.
//
// Operand forms:
//   K-major  (x W^T forward, dy W^T' input gradient): smem tile = 128 rows x 128 B, SWIZZLE_128B.
//   MN-major (dy^T x weight gradient, contraction over the batch): four 32(mn) x 32(k) boxes,
//            SWIZZLE_128B with 32-byte atoms -- the only MN-major layout tcgen05 accepts for tf32.
#include <cuda.h>

#include <unordered_map>

#include "gemm.cuh"

namespace porl {

constexpr int TC_BM = 128, TC_BN = 128, TC_BK = 32;      // tile; BK floats = one 128-byte swizzle span
constexpr int TC_STAGES = 3;
constexpr int TC_TILE_BYTES = TC_BM * TC_BK * 4;          // 16 KB per plane tile
constexpr int TC_STAGE_BYTES = 4 * TC_TILE_BYTES;         // A_hi, A_lo, B_hi, B_lo
constexpr int TC_EPI_FLOATS = 4 * 32 * 33;                // per-warp 32x33 transpose scratch
constexpr int TC_SMEM_BYTES = 1024 /*align slack*/ + TC_STAGES * TC_STAGE_BYTES + TC_EPI_FLOATS * 4 + 256;
constexpr int TC_THREADS = 192;
constexpr int TC_TMEM_COLS = 128;

struct alignas(64) TcParams {
  CUtensorMap mapA[kMaxGroups];
  CUtensorMap mapB[kMaxGroups];
  GemmBatch batch;
};

// ------------------------------------------------------------------------------------------ PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(bar), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                            int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout): start address, leading /
// stride byte offsets in 16-byte units, version 1 (sm_100), layout type in bits [61,64).
template <bool KMAJOR>
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr) {
  // K-major SWIZZLE_128B (type 2): 8-row groups of 128-byte rows are 1024 B apart (SBO); LBO unused.
  // MN-major SWIZZLE_128B_BASE32B (type 1): 4-row k-atoms 512 B apart (SBO); 32-wide mn blocks one
  // 32x32 box = 4096 B apart (LBO).
  constexpr uint64_t lbo = KMAJOR ? 1 : (4096 >> 4);
  constexpr uint64_t sbo = KMAJOR ? (1024 >> 4) : (512 >> 4);
  constexpr uint64_t type = KMAJOR ? 2 : 1;
  return (uint64_t)((addr >> 4) & 0x3fffu) | (lbo << 16) | (sbo << 32) | (1ull << 46) | (type << 61);
}

// ------------------------------------------------------------------------------------------ kernel
template <bool AK, bool BK>
__global__ void __launch_bounds__(TC_THREADS, 1) gemm_tc_kernel(const __grid_constant__ TcParams P) {
  const int gi = blockIdx.z % P.batch.n_groups;
  const int split = blockIdx.z / P.batch.n_groups;
  const GemmGroup& G = P.batch.g[gi];
  const int m0 = blockIdx.y * TC_BM, n0 = blockIdx.x * TC_BN;
  if (m0 >= G.M || n0 >= G.n_store) return;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;                         // SWIZZLE_128B tiles: 1024-B aligned
  uint8_t* base_ptr = smem_raw + (base - raw);
  float* epi = reinterpret_cast<float*>(base_ptr + TC_STAGES * TC_STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(base_ptr + TC_STAGES * TC_STAGE_BYTES + TC_EPI_FLOATS * 4);
  // bars[0..S) full, bars[S..2S) empty, bars[2S] tmem_full; then the TMEM base address word
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * TC_STAGES + 1);
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (TC_STAGES + s); };
  const uint32_t tmem_full_bar = bar0 + 8u * (2 * TC_STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = (G.K + TC_BK - 1) / TC_BK;
  const int kb0 = (int)((long long)split * nkb / P.batch.splits);
  const int kb1 = (int)((long long)(split + 1) * nkb / P.batch.splits);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&P.mapA[gi])) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&P.mapB[gi])) : "memory");
    for (int s = 0; s < TC_STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(TC_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(tmem_slot);

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int it = 0;
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const int s = it % TC_STAGES;
        const uint32_t ph = (uint32_t)(it / TC_STAGES) & 1u;
        mbar_wait(empty_bar(s), ph ^ 1u);
        mbar_expect_tx(full_bar(s), TC_STAGE_BYTES);
        const uint32_t st = base + (uint32_t)s * TC_STAGE_BYTES;
        const int k = kb * TC_BK;
#pragma unroll
        for (int pl = 0; pl < 2; ++pl) {
          const uint32_t a_dst = st + (uint32_t)pl * TC_TILE_BYTES;
          const uint32_t b_dst = st + (uint32_t)(2 + pl) * TC_TILE_BYTES;
          if (AK) {
            tma_load_3d(a_dst, &P.mapA[gi], full_bar(s), k, m0, pl);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) tma_load_3d(a_dst + j * 4096u, &P.mapA[gi], full_bar(s), m0 + 32 * j, k, pl);
          }
          if (BK) {
            tma_load_3d(b_dst, &P.mapB[gi], full_bar(s), k, n0, pl);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) tma_load_3d(b_dst + j * 4096u, &P.mapB[gi], full_bar(s), n0 + 32 * j, k, pl);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      // instruction descriptor: D fp32, A/B tf32, majors, N >> 3 at [17,23), M >> 4 at [24,29)
      constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((AK ? 0u : 1u) << 15) |
                                 ((BK ? 0u : 1u) << 16) | ((uint32_t)(TC_BN >> 3) << 17) |
                                 ((uint32_t)(TC_BM >> 4) << 24);
      constexpr uint32_t a_step = AK ? (32 >> 4) : (1024 >> 4);   // descriptor advance per 8-wide k-step
      constexpr uint32_t b_step = BK ? (32 >> 4) : (1024 >> 4);
      int it = 0;
      uint32_t acc = 0;
      for (int kb = kb0; kb < kb1; ++kb, ++it) {
        const int s = it % TC_STAGES;
        const uint32_t ph = (uint32_t)(it / TC_STAGES) & 1u;
        mbar_wait(full_bar(s), ph);
        tc_fence_after();
        const uint32_t st = base + (uint32_t)s * TC_STAGE_BYTES;
        const uint64_t a_hi = smem_desc<AK>(st), a_lo = smem_desc<AK>(st + TC_TILE_BYTES);
        const uint64_t b_hi = smem_desc<BK>(st + 2 * TC_TILE_BYTES), b_lo = smem_desc<BK>(st + 3 * TC_TILE_BYTES);
#pragma unroll
        for (int ks = 0; ks < TC_BK / 8; ++ks) {
          const uint64_t ao = (uint64_t)(a_step * ks), bo = (uint64_t)(b_step * ks);
          tc_mma_tf32(tmem_base, a_lo + ao, b_hi + bo, idesc, acc);   // small terms first
          acc = 1;
          tc_mma_tf32(tmem_base, a_hi + ao, b_lo + bo, idesc, 1u);
          tc_mma_tf32(tmem_base, a_hi + ao, b_hi + bo, idesc, 1u);
        }
        tc_commit(empty_bar(s));            // frees the stage once these MMAs have read it
      }
      tc_commit(tmem_full_bar);             // accumulator complete
    }
  } else {
    // ===================== epilogue: TMEM -> registers -> smem transpose -> global =====================
    const int q = warp & 3;                 // TMEM lane quarter this warp may read
    float* sc = epi + (warp - 2) * (32 * 33);
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    int* turn = nullptr;
    if (P.batch.splits > 1) {
      turn = P.batch.sem + gi * P.batch.sem_stride + blockIdx.y * gridDim.x + blockIdx.x;
      if (lane == 0) {
        while (ld_acquire(turn) != split) {
        }
      }
      __syncwarp();
    }
    const bool first = (split == 0);
    const int row_base = m0 + 32 * q;
    for (int c0 = 0; c0 < TC_BN; c0 += 32) {
      if (n0 + c0 >= G.n_store) break;
      float v[32];
      tc_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)c0, v);   // lane = row, 32 columns
#pragma unroll
      for (int j = 0; j < 32; ++j) sc[lane * 33 + j] = v[j];
      __syncwarp();
      const int col = n0 + c0 + lane;                                       // lane = column from here on
      const bool col_ok = col < G.n_store;
      const bool col_real = col < G.N;
      const float bias = (first && G.bias && col_real) ? G.bias[col] : 0.f;
      for (int r = 0; r < 32; ++r) {
        const int row = row_base + r;
        if (row >= G.M) break;
        float x = sc[r * 33 + lane];
        if (col_ok) {
          float* cp = G.C + (size_t)row * G.ldc + col;
          if (col_real) {
            if (!first) x += __ldcg(cp);
            x += bias;
            if (G.epilogue == 1) x = fmaxf(x, 0.f);
            else if (G.epilogue == 2) x = tanhf(x);
            if (G.relu_src && !(G.relu_src[(size_t)row * G.ld_relu + col] > 0.f)) x = 0.f;
          } else {
            x = 0.f;
          }
          if (G.C_lo) {
            const float h = tf32_hi(x);
            G.C_lo[(size_t)row * G.ldc + col] = tf32_hi(x - h);
            x = h;
          }
          *cp = x;
        }
      }
      __syncwarp();
    }
    if (P.batch.splits > 1) {
      __threadfence();
      // the four epilogue warps of this CTA release the turnstile together
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) st_release(turn, split + 1 == P.batch.splits ? 0 : split + 1);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TC_TMEM_COLS)
                 : "memory");
  }
}

// ------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

struct MapKey {
  const void* ptr;
  long long plane;   // lo - hi in floats
  int outer, inner, ld, kmajor;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && plane == o.plane && outer == o.outer && inner == o.inner && ld == o.ld && kmajor == o.kmajor;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.ptr);
    auto mix = [&](size_t v) { h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
    mix((size_t)k.plane); mix((size_t)k.outer); mix((size_t)k.inner); mix((size_t)k.ld); mix((size_t)k.kmajor);
    return h;
  }
};
static std::unordered_map<MapKey, CUtensorMap, MapKeyHash>& map_cache() {
  static thread_local std::unordered_map<MapKey, CUtensorMap, MapKeyHash> c;
  return c;
}

// 3-D tensor map over a (hi, lo) pair: dims {contiguous extent, strided extent, 2 planes}.
//   K-major operand  (rows x K, K contiguous):      box {32 k, 128 rows, 1}, SWIZZLE_128B
//   MN-major operand (K rows x MN, MN contiguous):  box {32 mn, 32 k, 1},    SWIZZLE_128B_ATOM_32B
static int make_map(CUtensorMap* out, const float* hi, const float* lo, int contiguous, int strided, int ld,
                    bool kmajor) {
  MapKey key{hi, (long long)(lo - hi), strided, contiguous, ld, kmajor ? 1 : 0};
  auto& cache = map_cache();
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return PORL_OK;
  }
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return PORL_ERR_CUDA;
  }
  const long long plane = lo - hi;
  if (plane <= 0 || (plane & 3)) return PORL_ERR_UNSUPPORTED;
  cuuint64_t dims[3] = {(cuuint64_t)contiguous, (cuuint64_t)strided, 2};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 4ull, (cuuint64_t)plane * 4ull};
  cuuint32_t box[3] = {32u, kmajor ? 128u : 32u, 1u};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(hi), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE,
                  kmajor ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) for extent %d x %d, ld %d", (int)r, strided, contiguous, ld);
    return PORL_ERR_CUDA;
  }
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *out);
  return PORL_OK;
}

template <bool AK, bool BK>
static int launch_cfg(const TcParams& P, dim3 grid, cudaStream_t st) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<AK, BK>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         TC_SMEM_BYTES);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(max dynamic smem %d) failed: %s", TC_SMEM_BYTES, cudaGetErrorString(e));
      return PORL_ERR_CUDA;
    }
    configured = true;
  }
  gemm_tc_kernel<AK, BK><<<grid, TC_THREADS, TC_SMEM_BYTES, st>>>(P);
  return PORL_OK;
}

int launch_gemm_tc(porl_ctx* ctx, bool a_kmajor, bool b_kmajor, GemmBatch& batch, cudaStream_t stream) {
  if (a_kmajor != b_kmajor) return PORL_ERR_UNSUPPORTED;     // forward / dx (both K-major) and dW (both MN-major)
  int max_m = 0, max_ns = 0, max_k = 0;
  long long tiles = 0;
  bool can_split = true;
  for (int i = 0; i < batch.n_groups; ++i) {
    const GemmGroup& g = batch.g[i];
    if (!g.A_lo || !g.B_lo) return PORL_ERR_UNSUPPORTED;      // operands must be (hi, lo) pairs
    if (g.colsum) return PORL_ERR_UNSUPPORTED;                 // bias gradients come from colsum_kernel
    if (g.N < 64 || g.K < 64 || g.M < 32) return PORL_ERR_UNSUPPORTED;
    if ((g.ldc & 3) || !aligned16(g.C)) return PORL_ERR_UNSUPPORTED;
    max_m = g.M > max_m ? g.M : max_m;
    max_ns = g.n_store > max_ns ? g.n_store : max_ns;
    max_k = g.K > max_k ? g.K : max_k;
    tiles += (long long)((g.M + TC_BM - 1) / TC_BM) * ((g.n_store + TC_BN - 1) / TC_BN);
    if (g.epilogue != 0 || g.relu_src != nullptr || g.C_lo != nullptr) can_split = false;
  }
  TcParams P;
  for (int i = 0; i < batch.n_groups; ++i) {
    const GemmGroup& g = batch.g[i];
    int rc;
    if (a_kmajor) {
      rc = make_map(&P.mapA[i], g.A, g.A_lo, g.K, g.M, g.lda, true);
      if (rc == PORL_OK) rc = make_map(&P.mapB[i], g.B, g.B_lo, g.K, g.N, g.ldb, true);
    } else {
      rc = make_map(&P.mapA[i], g.A, g.A_lo, g.M, g.K, g.lda, false);
      if (rc == PORL_OK) rc = make_map(&P.mapB[i], g.B, g.B_lo, g.N, g.K, g.ldb, false);
    }
    if (rc != PORL_OK) return rc;
  }
  dim3 grid((max_ns + TC_BN - 1) / TC_BN, (max_m + TC_BM - 1) / TC_BM, 1);
  int splits = 1;
  const int sms = ctx->sm_count;
  if (can_split && tiles < sms) {
    const int nkb = (max_k + TC_BK - 1) / TC_BK;
    int want = (int)((sms + tiles - 1) / tiles);
    int cap = nkb / 8;                       // at least 8 k-blocks (256 k) per slice
    splits = want < cap ? want : cap;
    if (splits < 1) splits = 1;
    if (splits > 16) splits = 16;
    int min_nkb = nkb;
    for (int i = 0; i < batch.n_groups; ++i) {
      const int n = (batch.g[i].K + TC_BK - 1) / TC_BK;
      min_nkb = n < min_nkb ? n : min_nkb;
    }
    if (splits > min_nkb) splits = min_nkb;  // never an empty k-slice (the accumulator would be undefined)
  }
  const int tiles_per_group = grid.x * grid.y;
  if (splits > 1 && (long long)tiles_per_group * batch.n_groups > kSemInts) splits = 1;
  batch.splits = splits;
  batch.sem = ctx->sem;
  batch.sem_stride = tiles_per_group;
  grid.z = batch.n_groups * splits;
  if (grid.y > 65535 || grid.z > 65535) return PORL_ERR_UNSUPPORTED;
  P.batch = batch;
  int rc = a_kmajor ? launch_cfg<true, true>(P, grid, stream) : launch_cfg<false, false>(P, grid, stream);
  if (rc != PORL_OK) return rc;
  ctx->tc_launches++;
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

}  // namespace porl
