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


// ---------------------------------------------------------------- fused scalar-V head step
// grid (row chunks, 2 online nets), 8 warps; a warp takes one row at a time (lanes stride the columns in float4s,
// NJ = ceil(H / 128) of them per lane): the two target rows and the online row are loaded together, three dot
// products give v1', v2', v_k, the TD target / expectile weight / dL/dv follow in registers (the formulas of
// expectile_kernel), and the same registers feed the head's backward: dz in place over h, dW and the bias gradient of
// the layer below as per-lane accumulators.  Head weights sit in shared memory.  Reductions over the batch are
// order-fixed: warps of a block in index order, row chunks in index order by the last block of the net (ticket), the
// loss by the last block of the grid.  Replaces head_forward (4 nets) + expectile_kernel + head_backward (2 nets): three
// launches and a second pass over the online activations.
constexpr int kVhGroup = 16;   // fan-in of the reduction tree
template <int NJ>
__global__ void __launch_bounds__(256) vhead_fused_kernel(const __grid_constant__ VHeadArgs a, float* partials,
                                                         uint32_t* tickets, uint32_t* status) {
  pdl_entry();
  extern __shared__ __align__(16) float vsm[];
  constexpr int W128 = NJ * 128;
  float* s_w = vsm;                       // [3][W128]: online w_k, target w_1, target w_2
  float* s_acc = vsm + 3 * W128;          // [8 warps][2][W128]: dW and db_prev accumulators
  __shared__ float s_scal[8][2];          // per warp: sum dv, loss
  __shared__ int s_flag[1];
  const int k = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = gridDim.x;
  const int Hp = (a.H + 3) & ~3;
  for (int i = threadIdx.x; i < 3 * W128; i += 256) {
    const int v = i / W128, c = i % W128;
    const float* src = v == 0 ? a.w[k] : a.wt[v - 1];
    s_w[i] = c < a.H ? __ldg(src + c) : 0.f;
  }
  __syncthreads();
  const float bk = a.b[k][0], bt1 = a.bt[0][0], bt2 = a.bt[1][0];
  float* const hk = a.h[k];
  const int64_t r0 = (int64_t)blockIdx.x * a.B / R, r1 = (int64_t)(blockIdx.x + 1) * a.B / R;
  float4 dw[NJ], dbp[NJ];
#pragma unroll
  for (int j = 0; j < NJ; ++j) dw[j] = dbp[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  float dbn = 0.f, loss = 0.f;
  for (int64_t row = r0 + warp; row < r1; row += 8) {
    float4 h4[NJ], t1[NJ], t2[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int c = 128 * j + 4 * lane;
      const bool ok = c < Hp;
      h4[j] = ok ? *reinterpret_cast<const float4*>(hk + row * a.ldh + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      t1[j] = ok ? *reinterpret_cast<const float4*>(a.ht[0] + row * a.ldh + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      t2[j] = ok ? *reinterpret_cast<const float4*>(a.ht[1] + row * a.ldh + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float s = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int c = 128 * j + 4 * lane;
      const float4 w4 = *reinterpret_cast<const float4*>(s_w + c);
      const float4 u4 = *reinterpret_cast<const float4*>(s_w + W128 + c);
      const float4 x4 = *reinterpret_cast<const float4*>(s_w + 2 * W128 + c);
      s = fmaf(h4[j].x, w4.x, s); s = fmaf(h4[j].y, w4.y, s); s = fmaf(h4[j].z, w4.z, s); s = fmaf(h4[j].w, w4.w, s);
      s1 = fmaf(t1[j].x, u4.x, s1); s1 = fmaf(t1[j].y, u4.y, s1); s1 = fmaf(t1[j].z, u4.z, s1); s1 = fmaf(t1[j].w, u4.w, s1);
      s2 = fmaf(t2[j].x, x4.x, s2); s2 = fmaf(t2[j].y, x4.y, s2); s2 = fmaf(t2[j].z, x4.z, s2); s2 = fmaf(t2[j].w, x4.w, s2);
    }
    const float v = warp_sum(s) + bk;
    const float nv = fminf(warp_sum(s1) + bt1, warp_sum(s2) + bt2);
    const float t = a.rewards[row] + ((1.f - a.dones[row]) * a.gamma) * nv;      // TD target (agent/por.py:85)
    const float u = t - v;
    const float wgt = fabsf(a.tau - (u < 0.f ? 1.f : 0.f));                        // asymmetric_l2_loss (por.py:15-17)
    const float dv = -(wgt * u) * a.inv_bn;
    if (lane == 0) {
      loss += wgt * u * u;
      if (k == 0) a.target_v[row] = t;
    }
    dbn += dv;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int c = 128 * j + 4 * lane;
      if (c < Hp) {
        const float4 w4 = *reinterpret_cast<const float4*>(s_w + c);
        float4 g = make_float4(dv * w4.x, dv * w4.y, dv * w4.z, dv * w4.w);
        if (!(h4[j].x > 0.f)) g.x = 0.f;
        if (!(h4[j].y > 0.f)) g.y = 0.f;
        if (!(h4[j].z > 0.f)) g.z = 0.f;
        if (!(h4[j].w > 0.f)) g.w = 0.f;
        *reinterpret_cast<float4*>(hk + row * a.ldh + c) = g;
        dw[j].x = fmaf(dv, h4[j].x, dw[j].x); dw[j].y = fmaf(dv, h4[j].y, dw[j].y);
        dw[j].z = fmaf(dv, h4[j].z, dw[j].z); dw[j].w = fmaf(dv, h4[j].w, dw[j].w);
        dbp[j].x += g.x; dbp[j].y += g.y; dbp[j].z += g.z; dbp[j].w += g.w;
      }
    }
  }
  // block partial: the 8 warps in index order
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const int c = 128 * j + 4 * lane;
    *reinterpret_cast<float4*>(s_acc + ((size_t)warp * 2 + 0) * W128 + c) = dw[j];
    *reinterpret_cast<float4*>(s_acc + ((size_t)warp * 2 + 1) * W128 + c) = dbp[j];
  }
  if (lane == 0) { s_scal[warp][0] = dbn; s_scal[warp][1] = loss; }
  __syncthreads();
  // partial layout per (net, row chunk): [2][W128] then [dv sum, loss]
  const size_t per_chunk = (size_t)2 * W128 + 4;
  float* part = partials + ((size_t)k * R + blockIdx.x) * per_chunk;
  for (int i = threadIdx.x; i < 2 * W128; i += 256) {
    float t = 0.f;
#pragma unroll
    for (int wq = 0; wq < 8; ++wq) t += s_acc[(size_t)wq * 2 * W128 + i];
    part[i] = t;
  }
  if (threadIdx.x < 4) {                                       // [dv sum, loss, 0, 0]: the tree reads whole float4s
    float t = 0.f;
    if (threadIdx.x < 2) {
#pragma unroll
      for (int wq = 0; wq < 8; ++wq) t += s_scal[wq][threadIdx.x];
    }
    part[2 * W128 + threadIdx.x] = t;
  }
  // Two-level, order-fixed reduction tree (fan-in kVhGroup): the last block of each group of kVhGroup consecutive row
  // chunks sums the group's partials (chunk order) into a group partial, the last group of the net to finish sums the
  // group partials (group order) into dW / db / db_prev, the last net adds the two loss sums.  One block summing all
  // R partials would pull R x 4 KB through a single SM (measured: +30 us at R = 256).
  const int G = (R + kVhGroup - 1) / kVhGroup, grp = blockIdx.x / kVhGroup;
  const int g0 = grp * kVhGroup, g1 = min(g0 + kVhGroup, R);
  float* const gpart = partials + (size_t)2 * R * per_chunk + ((size_t)k * G + grp) * per_chunk;
  auto last_of = [&](uint32_t* tk, uint32_t n) {          // true in every thread of the n-th block to arrive
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint32_t t = atomicAdd(tk, 1u);
      const bool last = (t == n - 1);
      if (last) *tk = 0;
      s_flag[0] = last ? 1 : 0;
    }
    __syncthreads();
    const bool last = s_flag[0] != 0;
    __syncthreads();
    if (last) __threadfence();
    return last;
  };
  auto sum_partials = [&](const float* src, int n, float* dst, bool final_net) {
    // four consecutive elements of n partials per thread: all n 16-byte loads in flight at once (one round trip for the
    // whole 4 KB vector; element by element it was five), each element added over the partials in index order
    auto emit = [&](int i, float t) {
      if (!final_net) {
        dst[i] = t;
      } else if (i < W128) {
        if (i < a.ldw) a.dW[k][i] = i < a.H ? t : 0.f;
      } else if (i < 2 * W128) {
        if (i - W128 < a.H) a.db_prev[k][i - W128] = t;
      } else if (i == 2 * W128) {
        a.db[k][0] = t;
      } else if (i == 2 * W128 + 1) {
        dst[i] = t;                                            // the net's loss sum, for the last net to add
      }
    };
    for (int i4 = threadIdx.x; i4 < (2 * W128 + 4) / 4; i4 += 256) {
      float4 vals[kVhGroup];
#pragma unroll
      for (int r = 0; r < kVhGroup; ++r)
        vals[r] = r < n ? __ldcg(reinterpret_cast<const float4*>(src + (size_t)r * per_chunk) + i4) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 t = vals[0];
#pragma unroll
      for (int r = 1; r < kVhGroup; ++r)
        if (r < n) { t.x += vals[r].x; t.y += vals[r].y; t.z += vals[r].z; t.w += vals[r].w; }
      emit(4 * i4, t.x); emit(4 * i4 + 1, t.y); emit(4 * i4 + 2, t.z); emit(4 * i4 + 3, t.w);
    }
  };
  if (!last_of(tickets + 8 + k * 64 + grp, (uint32_t)(g1 - g0))) return;
  sum_partials(partials + ((size_t)k * R + g0) * per_chunk, g1 - g0, gpart, false);
  if (!last_of(tickets + k, (uint32_t)G)) return;
  float* const npart = partials + (size_t)2 * R * per_chunk + (size_t)2 * G * per_chunk + (size_t)k * per_chunk;
  for (int c0 = 0; c0 < G; c0 += kVhGroup) {                   // G <= kVhGroup here (R <= kVhGroup^2); kept general
    sum_partials(partials + (size_t)2 * R * per_chunk + ((size_t)k * G + c0) * per_chunk, min(kVhGroup, G - c0), npart,
                 true);
  }
  if (!last_of(tickets + 2, 2u)) return;
  if (threadIdx.x == 0) {
    const float* n0 = partials + (size_t)2 * R * per_chunk + (size_t)2 * G * per_chunk;
    const float l = (__ldcg(n0 + 2 * W128 + 1) + __ldcg(n0 + per_chunk + 2 * W128 + 1)) * a.inv_bn * 0.5f;
    a.loss_out[0] = l;
    if (!isfinite(l)) atomicOr(status, (uint32_t)PORL_FLAG_NONFINITE_LOSS);
  }
}

// ---------------------------------------------------------------- fused policy head step
// The policy half of the update in one pass over the last hidden activations (agent/sorl.py:96-111, agent/por.py:97-108
// for a narrow Gaussian head): the scalar heads of the two value nets (new weights, forward only) give the advantage
// weight, the D <= 2 mean outputs of the policy head give the weighted Gaussian NLL (the formulas of awr_nll_kernel),
// and the same registers feed the head's backward: dz in place over h, dW, db, the bias gradient of the layer below and
// the log_std gradient.  grid (row chunks), 8 warps, a warp takes one row at a time.  Replaces head_forward (3 nets) +
// awr_nll_kernel + head_backward: three launches and two more passes over the activations.  Reductions over the batch
// are order-fixed like the value head's (warps in index order, then the two-level tree over the row chunks).
constexpr float kPhLog2Pi = 1.8378770664093453f;
constexpr float kPhWeightCap = 100.f;   // EXP_ADV_MAX, agent/por.py:12, agent/sorl.py:12
template <int NJ, int D>
__global__ void __launch_bounds__(256) phead_fused_kernel(const __grid_constant__ PHeadArgs a, float* partials,
                                                         uint32_t* tickets, uint32_t* status) {
  pdl_entry();
  extern __shared__ __align__(16) float psm2[];
  constexpr int W128 = NJ * 128;
  constexpr int NV = D + 1;               // vectors reduced over the batch: dW rows, db_prev
  constexpr int NS = 2 * D + 1;           // scalars: db[D], dlog_std[D], loss
  float* s_w = psm2;                      // [2 + D][W128]: value w_1, value w_2, policy W rows
  float* s_acc = psm2 + (2 + D) * W128;   // [8 warps][NV][W128]
  __shared__ float s_scal[8][NS];
  __shared__ int s_flag[1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = gridDim.x;
  const int Hp = (a.H + 3) & ~3;
  for (int i = threadIdx.x; i < (2 + D) * W128; i += 256) {
    const int v = i / W128, c = i % W128;
    const float* src = v < 2 ? a.wv[v] : a.w + (size_t)(v - 2) * a.ldw;
    s_w[i] = c < a.H ? __ldg(src + c) : 0.f;
  }
  __syncthreads();
  const float bv1 = a.bv[0][0], bv2 = a.bv[1][0];
  float bp[D], ls[D], inv_std[D];
  float nll_const = 0.5f * (float)D * kPhLog2Pi;
#pragma unroll
  for (int d = 0; d < D; ++d) {
    bp[d] = a.b[d];
    ls[d] = fminf(fmaxf(a.log_std[d], -5.f), 2.f);
    inv_std[d] = expf(-ls[d]);
    nll_const += ls[d];
  }
  float* const hp = a.h;
  const int64_t r0 = (int64_t)blockIdx.x * a.B / R, r1 = (int64_t)(blockIdx.x + 1) * a.B / R;
  float4 dw[D][NJ], dbp[NJ];
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    dbp[j] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int d = 0; d < D; ++d) dw[d][j] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float dbn[D], dls[D], loss = 0.f;
#pragma unroll
  for (int d = 0; d < D; ++d) dbn[d] = dls[d] = 0.f;
  for (int64_t row = r0 + warp; row < r1; row += 8) {
    float4 h4[NJ], t1[NJ], t2[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int c = 128 * j + 4 * lane;
      const bool ok = c < Hp;
      h4[j] = ok ? *reinterpret_cast<const float4*>(hp + row * a.ldh + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      t1[j] = ok ? *reinterpret_cast<const float4*>(a.hv[0] + row * a.ldh + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      t2[j] = ok ? *reinterpret_cast<const float4*>(a.hv[1] + row * a.ldh + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float s1 = 0.f, s2 = 0.f, sm[D];
#pragma unroll
    for (int d = 0; d < D; ++d) sm[d] = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int c = 128 * j + 4 * lane;
      const float4 u4 = *reinterpret_cast<const float4*>(s_w + c);
      const float4 x4 = *reinterpret_cast<const float4*>(s_w + W128 + c);
      s1 = fmaf(t1[j].x, u4.x, s1); s1 = fmaf(t1[j].y, u4.y, s1); s1 = fmaf(t1[j].z, u4.z, s1); s1 = fmaf(t1[j].w, u4.w, s1);
      s2 = fmaf(t2[j].x, x4.x, s2); s2 = fmaf(t2[j].y, x4.y, s2); s2 = fmaf(t2[j].z, x4.z, s2); s2 = fmaf(t2[j].w, x4.w, s2);
#pragma unroll
      for (int d = 0; d < D; ++d) {
        const float4 w4 = *reinterpret_cast<const float4*>(s_w + (2 + d) * W128 + c);
        sm[d] = fmaf(h4[j].x, w4.x, sm[d]); sm[d] = fmaf(h4[j].y, w4.y, sm[d]);
        sm[d] = fmaf(h4[j].z, w4.z, sm[d]); sm[d] = fmaf(h4[j].w, w4.w, sm[d]);
      }
    }
    const float v = fminf(warp_sum(s1) + bv1, warp_sum(s2) + bv2);
    const float adv = a.target_v[row] - v;
    const float e = (a.mode == PORL_AWR_POR) ? adv / a.alpha : a.alpha * adv;
    const float w = fminf(expf(e), kPhWeightCap);
    const float wn = w * a.inv_bn;
    float dmu[D], sz2 = 0.f;
#pragma unroll
    for (int d = 0; d < D; ++d) {
      float mu = warp_sum(sm[d]) + bp[d];
      if (a.tanh_out) mu = tanhf(mu);
      const float z = (a.x[row * a.ld_x + d] - mu) * inv_std[d];
      sz2 = fmaf(z, z, sz2);
      dmu[d] = -wn * z * inv_std[d];
      if (a.tanh_out) dmu[d] *= (1.f - mu * mu);
      dbn[d] += dmu[d];
      dls[d] += wn * (1.f - z * z);
    }
    const float nll = 0.5f * sz2 + nll_const;
    if (lane == 0) {
      if (!(nll > 0.f)) atomicOr(status, (uint32_t)PORL_FLAG_NONPOSITIVE_NLL);
      loss += w * nll;
    }
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int c = 128 * j + 4 * lane;
      if (c < Hp) {
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int d = 0; d < D; ++d) {
          const float4 w4 = *reinterpret_cast<const float4*>(s_w + (2 + d) * W128 + c);
          g.x = fmaf(dmu[d], w4.x, g.x); g.y = fmaf(dmu[d], w4.y, g.y); g.z = fmaf(dmu[d], w4.z, g.z); g.w = fmaf(dmu[d], w4.w, g.w);
          dw[d][j].x = fmaf(dmu[d], h4[j].x, dw[d][j].x); dw[d][j].y = fmaf(dmu[d], h4[j].y, dw[d][j].y);
          dw[d][j].z = fmaf(dmu[d], h4[j].z, dw[d][j].z); dw[d][j].w = fmaf(dmu[d], h4[j].w, dw[d][j].w);
        }
        if (!(h4[j].x > 0.f)) g.x = 0.f;
        if (!(h4[j].y > 0.f)) g.y = 0.f;
        if (!(h4[j].z > 0.f)) g.z = 0.f;
        if (!(h4[j].w > 0.f)) g.w = 0.f;
        *reinterpret_cast<float4*>(hp + row * a.ldh + c) = g;
        dbp[j].x += g.x; dbp[j].y += g.y; dbp[j].z += g.z; dbp[j].w += g.w;
      }
    }
  }
  // block partial: the 8 warps in index order
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const int c = 128 * j + 4 * lane;
#pragma unroll
    for (int d = 0; d < D; ++d) *reinterpret_cast<float4*>(s_acc + ((size_t)warp * NV + d) * W128 + c) = dw[d][j];
    *reinterpret_cast<float4*>(s_acc + ((size_t)warp * NV + D) * W128 + c) = dbp[j];
  }
  if (lane == 0) {
#pragma unroll
    for (int d = 0; d < D; ++d) { s_scal[warp][d] = dbn[d]; s_scal[warp][D + d] = dls[d]; }
    s_scal[warp][2 * D] = loss;
  }
  __syncthreads();
  // partial layout per row chunk: [NV][W128] then NS scalars (padded to 8)
  constexpr size_t per_chunk = (size_t)NV * W128 + 8;
  static_assert(NS <= 8, "scalar slots");
  float* part = partials + (size_t)blockIdx.x * per_chunk;
  for (int i = threadIdx.x; i < NV * W128; i += 256) {
    float t = 0.f;
#pragma unroll
    for (int wq = 0; wq < 8; ++wq) t += s_acc[(size_t)wq * NV * W128 + i];
    part[i] = t;
  }
  if (threadIdx.x < 8) {                                       // NS scalars, the rest of the 8 slots zero (read as float4s)
    float t = 0.f;
    if (threadIdx.x < NS) {
#pragma unroll
      for (int wq = 0; wq < 8; ++wq) t += s_scal[wq][threadIdx.x];
    }
    part[NV * W128 + threadIdx.x] = t;
  }
  // two-level, order-fixed reduction tree over the row chunks (fan-in kVhGroup), as in vhead_fused_kernel
  const int G = (R + kVhGroup - 1) / kVhGroup, grp = blockIdx.x / kVhGroup;
  const int g0 = grp * kVhGroup, g1 = min(g0 + kVhGroup, R);
  float* const gpart = partials + (size_t)R * per_chunk + (size_t)grp * per_chunk;
  auto last_of = [&](uint32_t* tk, uint32_t n) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint32_t t = atomicAdd(tk, 1u);
      const bool last = (t == n - 1);
      if (last) *tk = 0;
      s_flag[0] = last ? 1 : 0;
    }
    __syncthreads();
    const bool last = s_flag[0] != 0;
    __syncthreads();
    if (last) __threadfence();
    return last;
  };
  auto sum_partials = [&](const float* src, int n, float* dst, bool final) {
    auto emit = [&](int i, float t) {
      if (i >= NV * W128 + NS) return;
      if (!final) {
        dst[i] = t;
      } else if (i < D * W128) {
        const int d = i / W128, c = i % W128;
        if (c < a.ldw) a.dW[(size_t)d * a.ldw + c] = c < a.H ? t : 0.f;
      } else if (i < NV * W128) {
        if (i - D * W128 < a.H) a.db_prev[i - D * W128] = t;
      } else {
        const int sidx = i - NV * W128;
        if (sidx < D) {
          a.db[sidx] = t;
        } else if (sidx < 2 * D) {
          const float raw = a.log_std[sidx - D];
          a.dlog_std[sidx - D] = (raw >= -5.f && raw <= 2.f) ? t : 0.f;
        } else {
          const float l = t * a.inv_bn;
          a.loss_out[0] = l;
          if (!isfinite(l)) atomicOr(status, (uint32_t)PORL_FLAG_NONFINITE_LOSS);
        }
      }
    };
    // four consecutive elements of n partials per thread, all n 16-byte loads in flight (see vhead_fused_kernel)
    for (int i4 = threadIdx.x; i4 < (int)(per_chunk / 4); i4 += 256) {
      float4 vals[kVhGroup];
#pragma unroll
      for (int r = 0; r < kVhGroup; ++r)
        vals[r] = r < n ? __ldcg(reinterpret_cast<const float4*>(src + (size_t)r * per_chunk) + i4) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 t = vals[0];
#pragma unroll
      for (int r = 1; r < kVhGroup; ++r)
        if (r < n) { t.x += vals[r].x; t.y += vals[r].y; t.z += vals[r].z; t.w += vals[r].w; }
      emit(4 * i4, t.x); emit(4 * i4 + 1, t.y); emit(4 * i4 + 2, t.z); emit(4 * i4 + 3, t.w);
    }
  };
  if (!last_of(tickets + 256 + grp, (uint32_t)(g1 - g0))) return;
  sum_partials(partials + (size_t)g0 * per_chunk, g1 - g0, gpart, false);
  if (!last_of(tickets + 4, (uint32_t)G)) return;
  sum_partials(partials + (size_t)R * per_chunk, G, nullptr, true);            // G <= kVhGroup (R <= kVhGroup^2)
}

bool phead_fused_supported(int H, int D) { return H >= 4 && H <= 1024 && (D == 1 || D == 2); }

int launch_phead_fused(porl_ctx* ctx, const PHeadArgs& a, cudaStream_t st) {
  PORL_CHECK_ARG(phead_fused_supported(a.H, a.D) && (a.ldh & 3) == 0 && a.ldh >= ((a.H + 3) & ~3) && a.ldw >= a.H && a.B > 0,
                 "fused policy head: bad geometry (H=%d D=%d ldh=%d ldw=%d)", a.H, a.D, a.ldh, a.ldw);
  PORL_CHECK_ARG(a.hv[0] && a.hv[1] && a.wv[0] && a.wv[1] && a.bv[0] && a.bv[1] && a.h && a.w && a.b && a.log_std && a.x &&
                     a.target_v && a.dW && a.db && a.db_prev && a.dlog_std && a.loss_out && aligned16(a.h) &&
                     aligned16(a.hv[0]) && aligned16(a.hv[1]) && a.ld_x >= a.D && a.alpha != 0.f,
                 "fused policy head: null or misaligned pointer");
  const int nj = a.H <= 128 ? 1 : a.H <= 256 ? 2 : a.H <= 512 ? 4 : 8;
  const int w128 = nj * 128;
  PORL_CHECK_ARG(a.ldw <= w128, "fused policy head: ldw=%d exceeds %d", a.ldw, w128);
  int R = (int)((a.B + 15) / 16);
  int cap = ctx->sm_count;                 // one block per SM is resident (156 registers): one wave
  if (cap > kVhGroup * kVhGroup) cap = kVhGroup * kVhGroup;
  if (R > cap) R = cap;
  if (R < 1) R = 1;
  const size_t per_chunk = (size_t)(a.D + 1) * w128 + 8;
  while (R > 1 && ((size_t)R + (R + kVhGroup - 1) / kVhGroup + 1) * per_chunk > (size_t)kPartialFloats) R >>= 1;
  const size_t smem = (size_t)((2 + a.D) + 8 * (a.D + 1)) * w128 * sizeof(float);
  ProfScope prof(ctx, PORL_CAT_HEAD, 4.0 * (double)a.B * a.H * 4, st);
  uint32_t* tk = ctx->counters + kTicketBase;
  auto go = [&](auto kernel) -> int {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(fused policy head) failed: %s", cudaGetErrorString(e));
      return PORL_ERR_CUDA;
    }
    launch_pdl(ctx, kernel, dim3(R), 256, smem, st, a, ctx->partials, tk, ctx->counters);
    return PORL_OK;
  };
  int rc = PORL_ERR_UNSUPPORTED;
  if (a.D == 1) {
    rc = nj == 1 ? go(phead_fused_kernel<1, 1>) : nj == 2 ? go(phead_fused_kernel<2, 1>) : nj == 4 ? go(phead_fused_kernel<4, 1>)
                                                                                                    : go(phead_fused_kernel<8, 1>);
  } else {
    rc = nj == 1 ? go(phead_fused_kernel<1, 2>) : nj == 2 ? go(phead_fused_kernel<2, 2>) : nj == 4 ? go(phead_fused_kernel<4, 2>)
                                                                                                    : go(phead_fused_kernel<8, 2>);
  }
  if (rc != PORL_OK) return rc;
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

bool vhead_fused_supported(int H) { return H >= 4 && H <= 1024; }

int launch_vhead_fused(porl_ctx* ctx, const VHeadArgs& a, cudaStream_t st) {
  PORL_CHECK_ARG(vhead_fused_supported(a.H) && (a.ldh & 3) == 0 && a.ldh >= ((a.H + 3) & ~3) && a.ldw >= a.H && a.B > 0,
                 "fused value head: bad geometry (H=%d ldh=%d ldw=%d)", a.H, a.ldh, a.ldw);
  for (int k = 0; k < 2; ++k)
    PORL_CHECK_ARG(a.ht[k] && a.wt[k] && a.bt[k] && a.h[k] && a.w[k] && a.b[k] && a.dW[k] && a.db[k] && a.db_prev[k] &&
                       aligned16(a.ht[k]) && aligned16(a.h[k]),
                   "fused value head: null or misaligned pointer (net %d)", k);
  const int nj = a.H <= 128 ? 1 : a.H <= 256 ? 2 : a.H <= 512 ? 4 : 8;
  const int w128 = nj * 128;
  PORL_CHECK_ARG(a.ldw <= w128, "fused value head: ldw=%d exceeds %d", a.ldw, w128);
  // row chunks: at least 16 rows per block, and ONE wave of blocks -- grid (R, 2 nets) against the kernel's residency
  // (2 blocks per SM at H <= 512 by registers, 1 at H = 1024): a second wave repeats the weight staging and the rows'
  // load latency (ncu: 1.73 waves, `long scoreboard` the top stall)
  int R = (int)((a.B + 15) / 16);
  int cap = nj == 8 ? ctx->sm_count / 2 : ctx->sm_count;
  if (cap > kVhGroup * kVhGroup) cap = kVhGroup * kVhGroup;        // two levels of the reduction tree
  if (R > cap) R = cap;
  if (R < 1) R = 1;
  const size_t per_chunk = (size_t)2 * w128 + 4;
  while (R > 1 && (2 * (size_t)R + 2 * ((R + kVhGroup - 1) / kVhGroup) + 2) * per_chunk > (size_t)kPartialFloats) R >>= 1;
  const size_t smem = (size_t)(3 + 16) * w128 * sizeof(float);
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(vhead_fused_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 19 * 1024 * 4);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(vhead_fused_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 19 * 512 * 4);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(fused value head) failed: %s", cudaGetErrorString(e));
      return PORL_ERR_CUDA;
    }
    configured = true;
  }
  // bytes: three activations read, one written
  ProfScope prof(ctx, PORL_CAT_HEAD, 4.0 * (double)a.B * a.H * (2 * 3 + 2), st);
  uint32_t* tk = ctx->counters + kTicketBase;
  const dim3 grid(R, 2);
  switch (nj) {
    case 1: launch_pdl(ctx, vhead_fused_kernel<1>, grid, 256, smem, st, a, ctx->partials, tk, ctx->counters); break;
    case 2: launch_pdl(ctx, vhead_fused_kernel<2>, grid, 256, smem, st, a, ctx->partials, tk, ctx->counters); break;
    case 4: launch_pdl(ctx, vhead_fused_kernel<4>, grid, 256, smem, st, a, ctx->partials, tk, ctx->counters); break;
    default: launch_pdl(ctx, vhead_fused_kernel<8>, grid, 256, smem, st, a, ctx->partials, tk, ctx->counters); break;
  }
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

}  // namespace porl
