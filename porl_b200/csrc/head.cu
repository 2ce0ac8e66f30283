// Narrow output layers (scalar V heads, 2-D policy means, 5-way Q heads: n_out <= 8) as dedicated
// bandwidth kernels.  A GEMM tile is the wrong shape for them: y = h W^T with N = 1..8 is a row-wise
// dot product, and its backward is an outer product plus a reduction over the batch.  Both passes
// stream the (B x H) activation once, which is all the HBM/L2 traffic they need.
//
//   forward  : y[b, n] = act(sum_i h[b, i] W[n, i] + bias[n])                      (one warp per row)
//   backward : dW[n, i] = sum_b dout[b, n] h[b, i]        db[n] = sum_b dout[b, n]
//              dz[b, i] = 1[h[b, i] > 0] sum_n dout[b, n] W[n, i]   (in place over h)
//              db_prev[i] = sum_b dz[b, i]                 (bias gradient of the layer below, optional)
// Reductions over the batch are two-level and order-fixed (warp partials -> block partial -> the last
// block of a column chunk sums the row-chunk partials in index order): bitwise reproducible.
// Replaces the head Linear of util/util.py:29-47 mlp() and its autograd (agent/value_functions.py:31-42,
// agent/policy.py:38-49, src/porl/net/q_network.py:14-20).
#include <math.h>

#include "gemm.cuh"
#include "head.cuh"

namespace porl {

// ---------------------------------------------------------------- forward
template <int RPW>   // rows a warp processes at once: 4 for large batches (loads in flight), 1 for small ones (more blocks)
__global__ void __launch_bounds__(256) head_forward_kernel(const __grid_constant__ HeadFwdBatch hb) {
  pdl_entry();
  const HeadFwdSeg& s = hb.seg[blockIdx.y];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n_out = s.n_out;
  const int Hp = (s.H + 3) & ~3;
  // a warp takes RPW rows at a time: RPW x (Hp / 128) independent 16-byte loads in flight per lane
  for (int64_t row0 = ((int64_t)blockIdx.x * 8 + warp) * RPW; row0 < hb.B; row0 += (int64_t)gridDim.x * 8 * RPW) {
    float acc[RPW][kHeadMax];
#pragma unroll
    for (int u = 0; u < RPW; ++u)
#pragma unroll
      for (int n = 0; n < kHeadMax; ++n) acc[u][n] = 0.f;
    for (int c = lane * 4; c < Hp; c += 128) {
      float4 h4[RPW];
#pragma unroll
      for (int u = 0; u < RPW; ++u)
        h4[u] = (row0 + u < hb.B) ? *reinterpret_cast<const float4*>(s.h + (row0 + u) * s.ldh + c)
                                  : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int n = 0; n < kHeadMax; ++n) {
        if (n < n_out) {
          const float4 w4 = __ldg(reinterpret_cast<const float4*>(s.W + (size_t)n * s.ldw + c));
#pragma unroll
          for (int u = 0; u < RPW; ++u) {
            acc[u][n] = fmaf(h4[u].x, w4.x, acc[u][n]);
            acc[u][n] = fmaf(h4[u].y, w4.y, acc[u][n]);
            acc[u][n] = fmaf(h4[u].z, w4.z, acc[u][n]);
            acc[u][n] = fmaf(h4[u].w, w4.w, acc[u][n]);
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < RPW; ++u) {
      const int64_t row = row0 + u;
      if (row >= hb.B) break;
      float mine = 0.f;                      // lane n keeps output n
#pragma unroll
      for (int n = 0; n < kHeadMax; ++n) {
        if (n < n_out) {
          const float t = warp_sum(acc[u][n]);
          if (lane == n) mine = t;
        }
      }
      if (lane < s.ld_out) {
        float y = 0.f;
        if (lane < n_out) {
          y = mine + (s.bias ? s.bias[lane] : 0.f);
          if (s.act == 1) y = fmaxf(y, 0.f);
          else if (s.act == 2) y = tanhf(y);
        }
        s.out[row * s.ld_out + lane] = y;
      }
    }
  }
}

int launch_head_forward(porl_ctx* ctx, const HeadFwdBatch& hb, cudaStream_t st) {
  if (hb.n_seg <= 0) return PORL_OK;
  double bytes = 0;
  for (int i = 0; i < hb.n_seg; ++i) {
    const HeadFwdSeg& s = hb.seg[i];
    PORL_CHECK_ARG(s.h && s.W && s.out && s.n_out >= 1 && s.n_out <= kHeadMax && s.ld_out >= s.n_out &&
                       s.ld_out <= 32 && (s.ldh & 3) == 0 && (s.ldw & 3) == 0 && s.ldh >= ((s.H + 3) & ~3) &&
                       s.ldw >= ((s.H + 3) & ~3) && aligned16(s.h) && aligned16(s.W),
                   "head forward segment %d: bad geometry (n_out=%d H=%d ldh=%d ldw=%d ld_out=%d)", i, s.n_out, s.H,
                   s.ldh, s.ldw, s.ld_out);
    bytes += 4.0 * hb.B * (s.H + s.ld_out);
  }
  const bool wide = hb.B >= 2048;
  int64_t bx = wide ? (hb.B + 31) / 32 : (hb.B + 7) / 8;
  const int64_t cap = (int64_t)ctx->sm_count * 8;
  if (bx > cap) bx = cap;
  ProfScope prof(ctx, PORL_CAT_HEAD, bytes, st);
  if (wide) launch_pdl(ctx, head_forward_kernel<4>, dim3((unsigned)bx, hb.n_seg), 256, 0, st, hb);
  else launch_pdl(ctx, head_forward_kernel<1>, dim3((unsigned)bx, hb.n_seg), 256, 0, st, hb);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

// ---------------------------------------------------------------- backward
// grid (column chunks of 128, row chunks, nets); a lane owns 4 consecutive columns, the 8 warps of a block
// stride the rows of its row chunk.
__global__ void __launch_bounds__(256) head_backward_kernel(const __grid_constant__ HeadBwdBatch hb, float* partials,
                                                           uint32_t* tickets) {
  pdl_entry();
  __shared__ __align__(16) float sm[8][kHeadMax + 1][128];
  __shared__ float smb[8][kHeadMax];
  __shared__ int last_flag;
  const HeadBwdSeg& s = hb.seg[blockIdx.z];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n_out = hb.n_out;
  const int Hp = (hb.H + 3) & ~3;
  const int c = blockIdx.x * 128 + lane * 4;
  const bool col_ok = c < Hp;
  const int R = gridDim.y;
  const int64_t r0 = (int64_t)blockIdx.y * hb.B / R, r1 = (int64_t)(blockIdx.y + 1) * hb.B / R;

  float4 w4[kHeadMax], dw4[kHeadMax];
  float dbn[kHeadMax];
  float4 dbp = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int n = 0; n < kHeadMax; ++n) {
    w4[n] = (col_ok && n < n_out) ? __ldg(reinterpret_cast<const float4*>(s.W + (size_t)n * hb.ldw + c))
                                  : make_float4(0.f, 0.f, 0.f, 0.f);
    dw4[n] = make_float4(0.f, 0.f, 0.f, 0.f);
    dbn[n] = 0.f;
  }
#pragma unroll 2
  for (int64_t row = r0 + warp; row < r1; row += 8) {
    float d[kHeadMax];
#pragma unroll
    for (int n = 0; n < kHeadMax; ++n) d[n] = (n < n_out) ? s.dout[row * hb.ld_out + n] : 0.f;   // broadcast loads
    if (col_ok) {
      float* hp = s.h + row * hb.ldh + c;
      const float4 h4 = *reinterpret_cast<const float4*>(hp);
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int n = 0; n < kHeadMax; ++n) {
        if (n < n_out) {
          g.x = fmaf(d[n], w4[n].x, g.x); g.y = fmaf(d[n], w4[n].y, g.y);
          g.z = fmaf(d[n], w4[n].z, g.z); g.w = fmaf(d[n], w4[n].w, g.w);
          dw4[n].x = fmaf(d[n], h4.x, dw4[n].x); dw4[n].y = fmaf(d[n], h4.y, dw4[n].y);
          dw4[n].z = fmaf(d[n], h4.z, dw4[n].z); dw4[n].w = fmaf(d[n], h4.w, dw4[n].w);
        }
      }
      if (hb.need_dz) {
        if (!(h4.x > 0.f)) g.x = 0.f;
        if (!(h4.y > 0.f)) g.y = 0.f;
        if (!(h4.z > 0.f)) g.z = 0.f;
        if (!(h4.w > 0.f)) g.w = 0.f;
        *reinterpret_cast<float4*>(hp) = g;
        dbp.x += g.x; dbp.y += g.y; dbp.z += g.z; dbp.w += g.w;
      }
    }
#pragma unroll
    for (int n = 0; n < kHeadMax; ++n) dbn[n] += d[n];
  }
  // block partial: the 8 warps in index order
#pragma unroll
  for (int n = 0; n < kHeadMax; ++n) {
    if (n < n_out) *reinterpret_cast<float4*>(&sm[warp][n][lane * 4]) = dw4[n];
  }
  *reinterpret_cast<float4*>(&sm[warp][kHeadMax][lane * 4]) = dbp;
  if (lane == 0) {
#pragma unroll
    for (int n = 0; n < kHeadMax; ++n) smb[warp][n] = dbn[n];
  }
  __syncthreads();
  // partial layout: [net][row chunk][vector v = 0..n_out (n_out = db_prev)][Hp128] then [n_out] bias partials
  const int width = gridDim.x * 128;
  const size_t per_chunk = (size_t)(n_out + 1) * width + kHeadMax;
  float* part = partials + ((size_t)blockIdx.z * R + blockIdx.y) * per_chunk;
  for (int idx = threadIdx.x; idx < (n_out + 1) * 128; idx += 256) {
    const int v = idx >> 7, j = idx & 127;
    const int sv = v < n_out ? v : kHeadMax;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sm[w][sv][j];
    part[(size_t)v * width + blockIdx.x * 128 + j] = t;
  }
  if (blockIdx.x == 0 && threadIdx.x < n_out) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += smb[w][threadIdx.x];
    part[(size_t)(n_out + 1) * width + threadIdx.x] = t;
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
  if (!last_flag) return;
  __threadfence();
  const float* base = partials + (size_t)blockIdx.z * R * per_chunk;
  for (int idx = threadIdx.x; idx < (n_out + 1) * 128; idx += 256) {
    const int v = idx >> 7, j = idx & 127;
    const int col = blockIdx.x * 128 + j;
    float t = 0.f;
    for (int r = 0; r < R; ++r) t += __ldcg(base + (size_t)r * per_chunk + (size_t)v * width + col);
    if (v < n_out) {
      if (col < hb.ldw) s.dW[(size_t)v * hb.ldw + col] = col < hb.H ? t : 0.f;
    } else if (s.db_prev && col < hb.H) {
      s.db_prev[col] = t;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < n_out) {
    float t = 0.f;
    for (int r = 0; r < R; ++r) t += __ldcg(base + (size_t)r * per_chunk + (size_t)(n_out + 1) * width + threadIdx.x);
    s.db[threadIdx.x] = t;
  }
}

int launch_head_backward(porl_ctx* ctx, const HeadBwdBatch& hb, cudaStream_t st) {
  if (hb.n_seg <= 0) return PORL_OK;
  PORL_CHECK_ARG(hb.n_out >= 1 && hb.n_out <= kHeadMax && hb.ld_out >= hb.n_out && (hb.ldh & 3) == 0 &&
                     (hb.ldw & 3) == 0 && hb.ldh >= ((hb.H + 3) & ~3) && hb.ldw >= ((hb.H + 3) & ~3) && hb.B > 0,
                 "head backward: bad geometry (n_out=%d H=%d ldh=%d ldw=%d ld_out=%d)", hb.n_out, hb.H, hb.ldh, hb.ldw,
                 hb.ld_out);
  for (int i = 0; i < hb.n_seg; ++i) {
    const HeadBwdSeg& s = hb.seg[i];
    PORL_CHECK_ARG(s.h && s.dout && s.W && s.dW && s.db && aligned16(s.h) && aligned16(s.W) && aligned16(s.dW),
                   "head backward segment %d: null or misaligned pointer", i);
  }
  // every column of the padded weight row is produced (zeros past H), so the grid covers ldw
  const int col_chunks = (hb.ldw + 127) / 128;
  int R = (int)((hb.B + 63) / 64);              // >= 64 rows per block
  if (R > kHeadRowChunks) R = kHeadRowChunks;
  if (R < 1) R = 1;
  const size_t per_chunk = (size_t)(hb.n_out + 1) * col_chunks * 128 + kHeadMax;
  while (R > 1 && (size_t)hb.n_seg * R * per_chunk > (size_t)kPartialFloats) R >>= 1;
  PORL_CHECK_ARG((size_t)hb.n_seg * R * per_chunk <= (size_t)kPartialFloats && hb.n_seg * col_chunks <= kTickets,
                 "head backward: workspace too small for %d nets of width %d", hb.n_seg, hb.ldw);
  ProfScope prof(ctx, PORL_CAT_HEAD, 4.0 * hb.n_seg * (double)hb.B * hb.H * (hb.need_dz ? 2.0 : 1.0), st);
  launch_pdl(ctx, head_backward_kernel, dim3(col_chunks, R, hb.n_seg), 256, 0, st, hb, ctx->partials, ctx->counters + kTicketBase);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

}  // namespace porl
