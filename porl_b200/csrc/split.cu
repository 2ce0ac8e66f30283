// Helpers of the 3xTF32 mode: fp32 -> (hi, lo) plane split, transposed weight split, and the
// batch-sum of a pair (bias gradients).  All L2/HBM-bound elementwise or reduction kernels.
#include <stdlib.h>

#include "gemm.cuh"

namespace porl {

// ---------------------------------------------------------------- flat split
// variant 0: hi = RN_tf32(x), lo = x - hi (full fp32; the tensor core truncates it)
// variant 1: hi = RN_tf32(x), lo = RN_tf32(x - hi)
// variant 2: hi = x as is (the tensor core reads its top 19 bits), lo = RN_tf32(x - trunc_tf32(x))
__device__ __forceinline__ void split1(float v, int variant, float& h, float& l) {
  if (variant == 2) {
    h = v;
    l = tf32_hi(v - __uint_as_float(__float_as_uint(v) & 0xffffe000u));
  } else {
    h = tf32_hi(v);
    l = v - h;
    if (variant == 1) l = tf32_hi(l);
  }
}

__global__ void __launch_bounds__(256) split_kernel(const __grid_constant__ SplitBatch sb, int variant) {
  const SplitSeg& s = sb.seg[blockIdx.y];
  const int64_t n4 = s.n >> 2;
  const float4* src = reinterpret_cast<const float4*>(s.src);
  float4* hi = reinterpret_cast<float4*>(s.hi);
  float4* lo = reinterpret_cast<float4*>(s.lo);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = src[i];
    float4 h, l;
    split1(v.x, variant, h.x, l.x);
    split1(v.y, variant, h.y, l.y);
    split1(v.z, variant, h.z, l.z);
    split1(v.w, variant, h.w, l.w);
    hi[i] = h;
    lo[i] = l;
  }
}

int launch_split(porl_ctx* ctx, const SplitBatch& sb, cudaStream_t st) {
  if (sb.n_seg <= 0) return PORL_OK;
  int64_t max_n = 0, total = 0;
  for (int i = 0; i < sb.n_seg; ++i) {
    const SplitSeg& s = sb.seg[i];
    PORL_CHECK_ARG(s.src && s.hi && s.lo && (s.n & 3) == 0 && aligned16(s.src) && aligned16(s.hi) && aligned16(s.lo),
                   "split segment %d: needs 16-byte aligned buffers and n %% 4 == 0 (n=%lld)", i, (long long)s.n);
    max_n = s.n > max_n ? s.n : max_n;
    total += s.n;
  }
  int64_t bx = (max_n / 4 + 255) / 256;
  const int64_t cap = (int64_t)ctx->sm_count * 8;
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  ProfScope prof(ctx, PORL_CAT_SPLIT, (double)total * 12.0, st);
  static int variant = -1;
  if (variant < 0) {
    const char* e = getenv("PORL_SPLIT_VARIANT");
    variant = e ? atoi(e) : 1;
  }
  split_kernel<<<dim3((unsigned)bx, sb.n_seg), 256, 0, st>>>(sb, variant);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

// ---------------------------------------------------------------- transposed split (weights)
// src (rows, ld_src) logical (rows x cols)  ->  hi/lo (cols, ld_dst) = src^T, pad columns rows..ld_dst-1 zeroed
__global__ void __launch_bounds__(256) transpose_split_kernel(const __grid_constant__ TransposeBatch tb) {
  __shared__ float tile[32][33];
  const TransposeSeg& s = tb.seg[blockIdx.z];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;       // c0: source column block, r0: source row block
  if (c0 >= s.cols || r0 >= s.ld_dst) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // 32 x 8
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int r = r0 + ty + j, c = c0 + tx;
    tile[ty + j][tx] = (r < s.rows && c < s.cols) ? s.src[(size_t)r * s.ld_src + c] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int c = c0 + ty + j, r = r0 + tx;                    // destination row = source column
    if (c < s.cols && r < s.ld_dst) {
      const float v = tile[tx][ty + j];
      const float h = tf32_hi(v);
      s.hi[(size_t)c * s.ld_dst + r] = h;
      s.lo[(size_t)c * s.ld_dst + r] = tf32_hi(v - h);
    }
  }
}

int launch_transpose_split(porl_ctx* ctx, const TransposeBatch& tb, cudaStream_t st) {
  if (tb.n_seg <= 0) return PORL_OK;
  int max_c = 0, max_r = 0;
  double bytes = 0;
  for (int i = 0; i < tb.n_seg; ++i) {
    const TransposeSeg& s = tb.seg[i];
    PORL_CHECK_ARG(s.src && s.hi && s.lo && s.rows > 0 && s.cols > 0 && s.ld_dst >= s.rows && s.ld_src >= s.cols,
                   "transpose segment %d: bad shape", i);
    max_c = s.cols > max_c ? s.cols : max_c;
    max_r = s.ld_dst > max_r ? s.ld_dst : max_r;
    bytes += 12.0 * s.rows * s.cols;
  }
  ProfScope prof(ctx, PORL_CAT_SPLIT, bytes, st);
  transpose_split_kernel<<<dim3((max_c + 31) / 32, (max_r + 31) / 32, tb.n_seg), 256, 0, st>>>(tb);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

// ---------------------------------------------------------------- batch sum of a pair: out[n] = sum_b (hi+lo)[b, n]
// grid (column chunks of 32, row chunks, segments); 8 warps stride the rows of a chunk; the last block
// of a column chunk (ticket) adds the row-chunk partials in index order -> deterministic.
__global__ void __launch_bounds__(256) colsum_pair_kernel(const __grid_constant__ ColsumBatch cb, float* partials,
                                                          uint32_t* tickets) {
  __shared__ float sm[8][33];
  __shared__ int last_flag;
  const ColsumSeg& s = cb.seg[blockIdx.z];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + lane;
  if (blockIdx.x * 32 >= s.n) return;
  const int R = gridDim.y;
  const int64_t r0 = (int64_t)blockIdx.y * s.rows / R, r1 = (int64_t)(blockIdx.y + 1) * s.rows / R;
  float acc = 0.f;
  if (col < s.n) {
    for (int64_t r = r0 + warp; r < r1; r += 8) {
      float v = s.hi[r * s.ld + col];
      if (s.lo) v += s.lo[r * s.ld + col];
      acc += v;
    }
  }
  sm[warp][lane] = acc;
  __syncthreads();
  float* part = partials + ((size_t)blockIdx.z * R + blockIdx.y) * cb.max_n;
  if (warp == 0 && col < s.n) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sm[w][lane];
    part[col] = t;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* tk = tickets + blockIdx.z * gridDim.x + blockIdx.x;
    const uint32_t t = atomicAdd(tk, 1u);
    const bool last = (t == (uint32_t)R - 1);
    if (last) *tk = 0;
    last_flag = last ? 1 : 0;
  }
  __syncthreads();
  if (last_flag && warp == 0 && col < s.n) {
    __threadfence();
    float t = 0.f;
    for (int r = 0; r < R; ++r) t += __ldcg(partials + ((size_t)blockIdx.z * R + r) * cb.max_n + col);
    s.out[col] = t;
  }
}

int launch_colsum_pair(porl_ctx* ctx, ColsumBatch& cb, cudaStream_t st) {
  if (cb.n_seg <= 0) return PORL_OK;
  int max_n = 0;
  int64_t max_rows = 0;
  double bytes = 0;
  for (int i = 0; i < cb.n_seg; ++i) {
    const ColsumSeg& s = cb.seg[i];
    PORL_CHECK_ARG(s.hi && s.out && s.n > 0 && s.rows > 0 && s.ld >= s.n, "colsum segment %d: bad shape", i);
    max_n = s.n > max_n ? s.n : max_n;
    max_rows = s.rows > max_rows ? s.rows : max_rows;
    bytes += (s.lo ? 8.0 : 4.0) * s.rows * s.n;
  }
  cb.max_n = max_n;
  const int col_chunks = (max_n + 31) / 32;
  int R = (int)((max_rows + 255) / 256);      // >= 256 rows per block
  if (R > 32) R = 32;
  if (R < 1) R = 1;
  PORL_CHECK_ARG((int64_t)cb.n_seg * R * max_n <= kPartialFloats && cb.n_seg * col_chunks <= kTickets,
                 "colsum: workspace too small for %d segments of width %d", cb.n_seg, max_n);
  ProfScope prof(ctx, PORL_CAT_LOSS, bytes, st);
  colsum_pair_kernel<<<dim3(col_chunks, R, cb.n_seg), 256, 0, st>>>(cb, ctx->partials, ctx->counters + kTicketBase);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

}  // namespace porl
