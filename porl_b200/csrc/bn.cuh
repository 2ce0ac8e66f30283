// Train-mode BatchNorm over an (rows x C) activation matrix, shared by the FasterNet and PointMLP encoders.
#pragma once
#include "common.cuh"

namespace porl {

// ------------------------------------------------------------------------------------ BatchNorm (train mode)
// Column statistics of X[rows, C] in two levels: every block reduces a slab of rows (fp32 running sums flushed
// into fp64 every 16 rows, fixed order), writes (sum, sum of squares) per column as fp64 partials, and the last
// block to finish (ticket) adds the partials in block order into totals[C][2].  bn_finalize_kernel then emits
//     scale = gamma / sqrt(var_biased + eps),  shift = beta - mean * scale
// plus the running-stat update  r = (1 - m) r + m stat  (unbiased variance), nn.BatchNorm2d training forward;
// between the two, a data-parallel run sums the totals over the ranks (porl_fasternet_t.stats_allreduce).
// block = T threads = (T / (C/4)) row lanes x (C/4) float4 column groups; C/4 must divide T
// (T = 192 for FasterNet's C in {96, 192, 384}; T = 256 for PointMLP's powers of two up to 1024).
constexpr int kBnThreads = 192;

template <int T>
static __global__ void __launch_bounds__(T) bn_stats_kernel(const float* __restrict__ x, int64_t rows, int C,
                                                              int64_t rows_per_block, double* __restrict__ partial,
                                                              uint32_t* __restrict__ ticket, double* __restrict__ totals) {
  extern __shared__ double sred[];                 // [lanes][C][2]
  __shared__ bool last;
  const int c4n = C >> 2, lanes = T / c4n;
  const int cg = threadIdx.x % c4n, lane = threadIdx.x / c4n;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
  double s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0};
  float fs[4] = {0, 0, 0, 0}, fq[4] = {0, 0, 0, 0};
  int n = 0;
  auto add = [&](const float4 v) {
    fs[0] += v.x; fs[1] += v.y; fs[2] += v.z; fs[3] += v.w;
    fq[0] = fmaf(v.x, v.x, fq[0]); fq[1] = fmaf(v.y, v.y, fq[1]);
    fq[2] = fmaf(v.z, v.z, fq[2]); fq[3] = fmaf(v.w, v.w, fq[3]);
  };
  const float* xc = x + 4 * cg;
  int64_t r = r0 + lane;
  // four rows in flight per thread (independent 16-byte loads), fp32 running sums flushed to fp64 every 16 rows
  for (; r + 3 * (int64_t)lanes < r1; r += 4 * (int64_t)lanes) {
    const float4 v0 = *reinterpret_cast<const float4*>(xc + r * C);
    const float4 v1 = *reinterpret_cast<const float4*>(xc + (r + lanes) * C);
    const float4 v2 = *reinterpret_cast<const float4*>(xc + (r + 2 * (int64_t)lanes) * C);
    const float4 v3 = *reinterpret_cast<const float4*>(xc + (r + 3 * (int64_t)lanes) * C);
    add(v0); add(v1); add(v2); add(v3);
    n += 4;
    if (n == 16) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { s[j] += fs[j]; q[j] += fq[j]; fs[j] = 0.f; fq[j] = 0.f; }
      n = 0;
    }
  }
  for (; r < r1; r += lanes) add(*reinterpret_cast<const float4*>(xc + r * C));
#pragma unroll
  for (int j = 0; j < 4; ++j) { s[j] += fs[j]; q[j] += fq[j]; }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    sred[((size_t)lane * C + 4 * cg + j) * 2 + 0] = s[j];
    sred[((size_t)lane * C + 4 * cg + j) * 2 + 1] = q[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += T) {
    double a = 0, b = 0;
    for (int l = 0; l < lanes; ++l) { a += sred[((size_t)l * C + c) * 2]; b += sred[((size_t)l * C + c) * 2 + 1]; }
    partial[((size_t)blockIdx.x * C + c) * 2] = a;
    partial[((size_t)blockIdx.x * C + c) * 2 + 1] = b;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  __threadfence();
  for (int c = threadIdx.x; c < C; c += T) {
    double a = 0, b = 0;
    for (unsigned g = 0; g < gridDim.x; ++g) {
      a += __ldcg(&partial[((size_t)g * C + c) * 2]);
      b += __ldcg(&partial[((size_t)g * C + c) * 2 + 1]);
    }
    totals[2 * c] = a;
    totals[2 * c + 1] = b;
  }
  if (threadIdx.x == 0) *ticket = 0u;
}

// totals (sum, sum of squares per channel over `rows` rows, possibly summed over ranks) -> scale / shift and the
// running-statistic update of nn.BatchNorm2d's training forward
static __global__ void bn_finalize_kernel(const double* __restrict__ totals, double rows, int C, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* __restrict__ run_mean, float* __restrict__ run_var,
                                   float eps, float momentum, float* __restrict__ scale_shift) {
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const double mean = totals[2 * c] / rows;
    double var = totals[2 * c + 1] / rows - mean * mean;
    if (var < 0) var = 0;
    const float sc = (float)((double)gamma[c] / sqrt(var + (double)eps));
    scale_shift[c] = sc;
    scale_shift[C + c] = (float)((double)beta[c] - mean * (double)sc);
    const double unbiased = rows > 1 ? var * rows / (rows - 1) : var;
    run_mean[c] = (float)((1.0 - (double)momentum) * (double)run_mean[c] + (double)momentum * mean);
    run_var[c] = (float)((1.0 - (double)momentum) * (double)run_var[c] + (double)momentum * unbiased);
  }
}

// y = x * scale[c] + shift[c] (+ ReLU), in place; float4 over an (rows x C) matrix
static __global__ void __launch_bounds__(256) bn_apply_kernel(float* __restrict__ x, int64_t n4, int c4n,
                                                       const float* __restrict__ scale_shift, int C, int relu) {
  const float4* sc4 = reinterpret_cast<const float4*>(scale_shift);
  const float4* sh4 = reinterpret_cast<const float4*>(scale_shift + C);
  float4* x4 = reinterpret_cast<float4*>(x);
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
    const int cg = (int)(i % c4n);
    const float4 sc = __ldg(sc4 + cg), sh = __ldg(sh4 + cg);
    float4 v = x4[i];
    v.x = fmaf(v.x, sc.x, sh.x); v.y = fmaf(v.y, sc.y, sh.y); v.z = fmaf(v.z, sc.z, sh.z); v.w = fmaf(v.w, sc.w, sh.w);
    if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
    x4[i] = v;
  }
}


}  // namespace porl
