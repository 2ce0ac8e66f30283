// fp32 FFMA grouped contraction ("exact" mode and the small-shape path).
//
// Classic register-tiled SGEMM: 256 threads, BK = 16, 128x128 (8x8 per thread) or 64x64 (4x4 per
// thread) tiles, double-buffered shared memory with register prefetch, XOR-swizzled tiles so that
// both the transposing stores of K-major operands and the float4 fragment loads are conflict-free.
// Split-K runs through per-tile turnstile words (a CTA adds its partial after the CTA of the
// previous K-slice of the same tile), which keeps the summation order fixed -> bitwise reproducible.
#include "gemm.cuh"

namespace porl {

constexpr int BKT = 16;

template <int BM, bool KMAJOR>
struct TileLoader {
  // A tile of BM "outer" elements x BKT k-elements -> 4*BM floats = BM float4; 256 threads.
  static constexpr int kVec = BM * BKT / 4 / 256;
  float4 v[kVec];

  // P: operand base; ld: leading dimension; o0: first outer index; outer_n: outer extent;
  // k0: first k; k_n: K extent.
  __device__ __forceinline__ void load(const float* __restrict__ P, const float* __restrict__ Plo, int ld,
                                       int o0, int outer_n, int k0, int k_n, int tid) {
#pragma unroll
    for (int i = 0; i < kVec; ++i) {
      const int f = tid + i * 256;
      float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
      if (KMAJOR) {
        const int row = f >> 2, kq = (f & 3) << 2;
        const int o = o0 + row, k = k0 + kq;
        if (o < outer_n && k < k_n) {
          r = *reinterpret_cast<const float4*>(P + (size_t)o * ld + k);
          if (Plo) {
            const float4 q = *reinterpret_cast<const float4*>(Plo + (size_t)o * ld + k);
            r.x += q.x; r.y += q.y; r.z += q.z; r.w += q.w;
          }
          if (k + 1 >= k_n) r.y = 0.f;
          if (k + 2 >= k_n) r.z = 0.f;
          if (k + 3 >= k_n) r.w = 0.f;
        }
      } else {
        const int kr = f / (BM / 4), c4 = (f % (BM / 4)) << 2;
        const int k = k0 + kr, o = o0 + c4;
        if (k < k_n && o < outer_n) {
          r = *reinterpret_cast<const float4*>(P + (size_t)k * ld + o);
          if (Plo) {
            const float4 q = *reinterpret_cast<const float4*>(Plo + (size_t)k * ld + o);
            r.x += q.x; r.y += q.y; r.z += q.z; r.w += q.w;
          }
          if (o + 1 >= outer_n) r.y = 0.f;
          if (o + 2 >= outer_n) r.z = 0.f;
          if (o + 3 >= outer_n) r.w = 0.f;
        }
      }
      v[i] = r;
    }
  }

  // tile layout: T[k][outer ^ swz(k)], swz(k) = ((k >> 2) & 3) << 3
  __device__ __forceinline__ void store(float (*T)[BM], int tid) const {
#pragma unroll
    for (int i = 0; i < kVec; ++i) {
      const int f = tid + i * 256;
      if (KMAJOR) {
        const int row = f >> 2, q = f & 3, kq = q << 2;
        const int col = row ^ (q << 3);
        T[kq + 0][col] = v[i].x;
        T[kq + 1][col] = v[i].y;
        T[kq + 2][col] = v[i].z;
        T[kq + 3][col] = v[i].w;
      } else {
        const int kr = f / (BM / 4), c4 = (f % (BM / 4)) << 2;
        const int col = c4 ^ (((kr >> 2) & 3) << 3);
        *reinterpret_cast<float4*>(&T[kr][col]) = v[i];
      }
    }
  }
};

template <int BM, int BN, bool AK, bool BK>
__global__ void __launch_bounds__(256, (BM == 128 ? 2 : 3))
gemm_simt_kernel(const __grid_constant__ GemmBatch batch) {
  constexpr int TM = BM / 16, TN = BN / 16;       // per-thread tile (16 x 16 threads)
  constexpr int CM = TM / 4, CN = TN / 4;         // float4 chunks per thread along M / N
  constexpr int SM_ = BM / CM, SN_ = BN / CN;     // distance between a thread's chunks

  const int gi = blockIdx.z % batch.n_groups;
  const int split = blockIdx.z / batch.n_groups;
  const GemmGroup& G = batch.g[gi];
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  if (m0 >= G.M || n0 >= G.n_store) return;

  __shared__ __align__(16) float As[2][BKT][BM];
  __shared__ __align__(16) float Bs[2][BKT][BN];

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;

  const int nkt = (G.K + BKT - 1) / BKT;
  const int kt_begin = (int)((long long)split * nkt / batch.splits);
  const int kt_end = (int)((long long)(split + 1) * nkt / batch.splits);

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
  float asum[TM];
#pragma unroll
  for (int i = 0; i < TM; ++i) asum[i] = 0.f;
  const bool do_colsum = (G.colsum != nullptr) && blockIdx.x == 0 && tx == 0;

  TileLoader<BM, AK> la;
  TileLoader<BN, BK> lb;

  if (kt_begin < kt_end) {
    la.load(G.A, G.A_lo, G.lda, m0, G.M, kt_begin * BKT, G.K, tid);
    lb.load(G.B, G.B_lo, G.ldb, n0, G.N, kt_begin * BKT, G.K, tid);
    la.store(As[0], tid);
    lb.store(Bs[0], tid);
  }
  __syncthreads();

  for (int kt = kt_begin; kt < kt_end; ++kt) {
    const int buf = (kt - kt_begin) & 1;
    const bool has_next = kt + 1 < kt_end;
    if (has_next) {
      la.load(G.A, G.A_lo, G.lda, m0, G.M, (kt + 1) * BKT, G.K, tid);
      lb.load(G.B, G.B_lo, G.ldb, n0, G.N, (kt + 1) * BKT, G.K, tid);
    }
#pragma unroll
    for (int kk = 0; kk < BKT; ++kk) {
      const int s = ((kk >> 2) & 3) << 3;
      float a[TM], b[TN];
#pragma unroll
      for (int c = 0; c < CM; ++c) {
        const float4 t = *reinterpret_cast<const float4*>(&As[buf][kk][(c * SM_ + ty * 4) ^ s]);
        a[c * 4 + 0] = t.x; a[c * 4 + 1] = t.y; a[c * 4 + 2] = t.z; a[c * 4 + 3] = t.w;
      }
#pragma unroll
      for (int c = 0; c < CN; ++c) {
        const float4 t = *reinterpret_cast<const float4*>(&Bs[buf][kk][(c * SN_ + tx * 4) ^ s]);
        b[c * 4 + 0] = t.x; b[c * 4 + 1] = t.y; b[c * 4 + 2] = t.z; b[c * 4 + 3] = t.w;
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      if (do_colsum) {
#pragma unroll
        for (int i = 0; i < TM; ++i) asum[i] += a[i];
      }
    }
    if (has_next) {
      la.store(As[buf ^ 1], tid);
      lb.store(Bs[buf ^ 1], tid);
    }
    __syncthreads();
  }

  // ---- epilogue (turnstile-ordered when split-K) ----
  int* turn = nullptr;
  if (batch.splits > 1) {
    turn = batch.sem + gi * batch.sem_stride + blockIdx.y * gridDim.x + blockIdx.x;
    if (tid == 0) {
      while (ld_acquire(turn) != split) {
      }
    }
    __syncthreads();
  }
  const bool first = (split == 0);
  const bool vec = ((G.ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(G.C) & 15u) == 0);

#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int row = m0 + (i >> 2) * SM_ + ty * 4 + (i & 3);
    if (row >= G.M) continue;
    float* crow = G.C + (size_t)row * G.ldc;
    const float* rrow = G.relu_src ? G.relu_src + (size_t)row * G.ld_relu : nullptr;
#pragma unroll
    for (int c = 0; c < CN; ++c) {
      const int col = n0 + c * SN_ + tx * 4;
      if (col >= G.n_store) continue;
      float o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float v = acc[i][c * 4 + j];
        const int cj = col + j;
        if (cj < G.N) {
          if (!first) v += __ldcg(crow + cj);
          if (first && G.bias) v += G.bias[cj];
          if (G.epilogue == 1) v = fmaxf(v, 0.f);
          else if (G.epilogue == 2) v = tanhf(v);
          if (rrow && !(rrow[cj] > 0.f)) v = 0.f;
        } else {
          v = 0.f;
        }
        o[j] = v;
      }
      if (G.C_lo) {   // pair output: hi = tf32(v), lo = v - hi
        float* lrow = G.C_lo + (size_t)row * G.ldc;
        float l[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float h = tf32_hi(o[j]);
          l[j] = tf32_hi(o[j] - h);
          o[j] = h;
        }
        if (vec && col + 3 < G.n_store) {
          *reinterpret_cast<float4*>(lrow + col) = make_float4(l[0], l[1], l[2], l[3]);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (col + j < G.n_store) lrow[col + j] = l[j];
        }
      }
      if (vec && col + 3 < G.n_store) {
        *reinterpret_cast<float4*>(crow + col) = make_float4(o[0], o[1], o[2], o[3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (col + j < G.n_store) crow[col + j] = o[j];
      }
    }
  }
  if (do_colsum) {
#pragma unroll
    for (int i = 0; i < TM; ++i) {
      const int row = m0 + (i >> 2) * SM_ + ty * 4 + (i & 3);
      if (row < G.M) G.colsum[row] = first ? asum[i] : __ldcg(G.colsum + row) + asum[i];
    }
  }
  if (batch.splits > 1) {
    __threadfence();
    __syncthreads();
    if (tid == 0) st_release(turn, split + 1 == batch.splits ? 0 : split + 1);
  }
}

template <int BM, int BN>
static void launch_cfg(bool ak, bool bk, const GemmBatch& b, dim3 grid, cudaStream_t st) {
  if (ak && bk) gemm_simt_kernel<BM, BN, true, true><<<grid, 256, 0, st>>>(b);
  else if (ak && !bk) gemm_simt_kernel<BM, BN, true, false><<<grid, 256, 0, st>>>(b);
  else if (!ak && bk) gemm_simt_kernel<BM, BN, false, true><<<grid, 256, 0, st>>>(b);
  else gemm_simt_kernel<BM, BN, false, false><<<grid, 256, 0, st>>>(b);
}

int launch_gemm_simt(porl_ctx* ctx, bool a_kmajor, bool b_kmajor, GemmBatch& batch, cudaStream_t stream) {
  int max_m = 0, max_ns = 0, max_k = 0;
  long long tiles128 = 0, tiles64 = 0;
  bool can_split = true;
  for (int i = 0; i < batch.n_groups; ++i) {
    const GemmGroup& g = batch.g[i];
    max_m = g.M > max_m ? g.M : max_m;
    max_ns = g.n_store > max_ns ? g.n_store : max_ns;
    max_k = g.K > max_k ? g.K : max_k;
    tiles128 += (long long)((g.M + 127) / 128) * ((g.n_store + 127) / 128);
    tiles64 += (long long)((g.M + 63) / 64) * ((g.n_store + 63) / 64);
    if (g.epilogue != 0 || g.relu_src != nullptr || g.C_lo != nullptr) can_split = false;
  }
  const int sms = ctx->sm_count;
  const bool big = tiles128 >= (long long)sms;   // enough 128x128 tiles for one CTA per SM
  const int BM = big ? 128 : 64;
  const long long tiles = big ? tiles128 : tiles64;
  dim3 grid((max_ns + BM - 1) / BM, (max_m + BM - 1) / BM, 1);
  int splits = 1;
  if (can_split && tiles < 2LL * sms) {
    const int nkt = (max_k + BKT - 1) / BKT;
    int want = (int)((2LL * sms + tiles - 1) / tiles);
    int cap = nkt / 8;   // at least 8 k-tiles (128 k) per slice
    splits = want < cap ? want : cap;
    if (splits < 1) splits = 1;
    if (splits > 32) splits = 32;
  }
  const int tiles_per_group = grid.x * grid.y;
  if (splits > 1 && (long long)tiles_per_group * batch.n_groups > kSemInts) splits = 1;
  batch.splits = splits;
  batch.sem = ctx->sem;
  batch.sem_stride = tiles_per_group;
  grid.z = batch.n_groups * splits;
  if (grid.y > 65535 || grid.z > 65535) {
    set_error("contraction grid too large (M=%d)", max_m);
    return PORL_ERR_INVALID;
  }
  if (big) launch_cfg<128, 128>(a_kmajor, b_kmajor, batch, grid, stream);
  else launch_cfg<64, 64>(a_kmajor, b_kmajor, batch, grid, stream);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

}  // namespace porl
