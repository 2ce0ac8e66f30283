// Helpers of the tensor-core mode: batched fp32 transposes (the tcgen05 kernel takes K-major operands
// only, so weight-gradient and input-gradient contractions read transposed copies) and the batch-sum
// that yields bias gradients.  L2/HBM-bound copy and reduction kernels.
#include "gemm.cuh"

namespace porl {

// ---------------------------------------------------------------- batched transpose
// dst (cols, ld_dst) = src (rows, ld_src)^T through a 32x33 shared-memory tile; both sides coalesced.
__global__ void __launch_bounds__(256) transpose_kernel(const __grid_constant__ TransposeBatch tb) {
  pdl_entry();
  __shared__ float tile[32][33];
  const TransposeSeg& s = tb.seg[blockIdx.z];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;       // c0: source column block, r0: source row block
  if (c0 >= s.cols || r0 >= s.rows) return;
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
    if (c < s.cols && r < s.rows) s.dst[(size_t)c * s.ld_dst + r] = tile[tx][ty + j];
  }
}

int launch_transpose(porl_ctx* ctx, const TransposeBatch& tb, cudaStream_t st) {
  if (tb.n_seg <= 0) return PORL_OK;
  int max_c = 0, max_r = 0;
  double bytes = 0;
  for (int i = 0; i < tb.n_seg; ++i) {
    const TransposeSeg& s = tb.seg[i];
    PORL_CHECK_ARG(s.src && s.dst && s.rows > 0 && s.cols > 0 && s.ld_dst >= s.rows && s.ld_src >= s.cols,
                   "transpose segment %d: bad shape (%d x %d, ld_src %d, ld_dst %d)", i, s.rows, s.cols, s.ld_src,
                   s.ld_dst);
    max_c = s.cols > max_c ? s.cols : max_c;
    max_r = s.rows > max_r ? s.rows : max_r;
    bytes += 8.0 * s.rows * s.cols;
  }
  PORL_CHECK_ARG((max_r + 31) / 32 <= 65535, "transpose: too many rows (%d)", max_r);
  ProfScope prof(ctx, PORL_CAT_TRANSPOSE, bytes, st);
  launch_pdl(ctx, transpose_kernel, dim3((max_c + 31) / 32, (max_r + 31) / 32, tb.n_seg), 256, 0, st, tb);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

// ---------------------------------------------------------------- batch sum: out[n] = sum_b src[b, n]
// grid (column chunks of 32, row chunks, segments); 8 warps stride the rows of a chunk; the last block
// of a column chunk (ticket) adds the row-chunk partials in index order -> deterministic.
__global__ void __launch_bounds__(256) colsum_kernel(const __grid_constant__ ColsumBatch cb, float* partials,
                                                     uint32_t* tickets) {
  pdl_entry();
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
    // four independent loads in flight per lane (a dependent acc += load chain ran at DRAM latency)
    int64_t r = r0 + warp;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    for (; r + 24 < r1; r += 32) {
      const float v0 = s.src[r * s.ld + col], v1 = s.src[(r + 8) * s.ld + col];
      const float v2 = s.src[(r + 16) * s.ld + col], v3 = s.src[(r + 24) * s.ld + col];
      a0 += v0; a1 += v1; a2 += v2; a3 += v3;
    }
    for (; r < r1; r += 8) a0 += s.src[r * s.ld + col];
    acc = (a0 + a1) + (a2 + a3);
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
    // all R <= 32 partials in flight at once (one round trip), added in row-chunk order
    float vals[32];
#pragma unroll
    for (int r = 0; r < 32; ++r) vals[r] = r < R ? __ldcg(partials + ((size_t)blockIdx.z * R + r) * cb.max_n + col) : 0.f;
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 32; ++r)
      if (r < R) t += vals[r];
    s.out[col] = t;
  }
}

int launch_colsum(porl_ctx* ctx, ColsumBatch& cb, cudaStream_t st) {
  if (cb.n_seg <= 0) return PORL_OK;
  int max_n = 0;
  int64_t max_rows = 0;
  double bytes = 0;
  for (int i = 0; i < cb.n_seg; ++i) {
    const ColsumSeg& s = cb.seg[i];
    PORL_CHECK_ARG(s.src && s.out && s.n > 0 && s.rows > 0 && s.ld >= s.n, "colsum segment %d: bad shape", i);
    max_n = s.n > max_n ? s.n : max_n;
    max_rows = s.rows > max_rows ? s.rows : max_rows;
    bytes += 4.0 * s.rows * s.n;
  }
  cb.max_n = max_n;
  const int col_chunks = (max_n + 31) / 32;
  int R = (int)((max_rows + 127) / 128);      // >= 128 rows per block
  if (R > 32) R = 32;
  if (R < 1) R = 1;
  PORL_CHECK_ARG((int64_t)cb.n_seg * R * max_n <= kPartialFloats && cb.n_seg * col_chunks <= kTickets,
                 "colsum: workspace too small for %d segments of width %d", cb.n_seg, max_n);
  ProfScope prof(ctx, PORL_CAT_LOSS, bytes, st);
  launch_pdl(ctx, colsum_kernel, dim3(col_chunks, R, cb.n_seg), 256, 0, st, cb, ctx->partials, ctx->counters + kTicketBase);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

}  // namespace porl

using namespace porl;

extern "C" int porl_transpose(porl_ctx_t* ctx, const float* src, int64_t rows, int64_t cols, int64_t ld_src,
                              float* dst, int64_t ld_dst, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && rows > 0 && cols > 0 && rows < (1 << 30) && cols < (1 << 30), "porl_transpose: bad shape");
  TransposeBatch tb;
  tb.n_seg = 1;
  tb.seg[0] = TransposeSeg{src, dst, (int)rows, (int)cols, (int)ld_src, (int)ld_dst};
  return launch_transpose(ctx, tb, as_stream(stream));
}
