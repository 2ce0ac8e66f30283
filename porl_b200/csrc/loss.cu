// K4 / K5 / K6: loss epilogues with their analytic gradients.
//
// Every scalar reduction is two-level and order-fixed: a block reduces its samples, writes one
// partial, and the last block to finish (atomic ticket) sums the partials in index order, so the
// losses and the log_std gradient are bitwise reproducible run to run.
#include <math.h>

#include <mutex>

#include "common.cuh"

namespace porl {

constexpr float kLog2Pi = 1.8378770664093453f;
constexpr float kWeightCap = 100.f;  // EXP_ADV_MAX, agent/por.py:12, agent/sorl.py:12

// returns true in every thread of the LAST block to arrive
__device__ __forceinline__ bool last_block_ticket(uint32_t* counter, int* smem_flag) {
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t t = atomicAdd(counter, 1u);
    const bool last = (t == gridDim.x - 1);
    if (last) *counter = 0;  // self-reset for the next launch (stream-ordered)
    *smem_flag = last ? 1 : 0;
  }
  __syncthreads();
  const bool last = *smem_flag != 0;
  if (last) __threadfence();
  return last;
}

// ---------------------------------------------------------------- K4
__global__ void __launch_bounds__(256)
expectile_kernel(const float* __restrict__ v1, const float* __restrict__ v2,
                 const float* __restrict__ nv1, const float* __restrict__ nv2, int64_t ldv,
                 const float* __restrict__ rewards, const float* __restrict__ dones, float gamma,
                 float tau, int64_t B, float inv_bn, float* __restrict__ target_v,
                 float* __restrict__ dv1, float* __restrict__ dv2, float* __restrict__ loss_out,
                 float* partials, uint32_t* counter, uint32_t* status) {
  pdl_entry();
  __shared__ float red[32];
  __shared__ int flag;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float l = 0.f;
  if (i < B) {
    const float nv = fminf(nv1[i * ldv], nv2[i * ldv]);
    const float t = rewards[i] + ((1.f - dones[i]) * gamma) * nv;
    target_v[i] = t;
    {
      const float u = t - v1[i * ldv];
      const float w = fabsf(tau - (u < 0.f ? 1.f : 0.f));
      l += w * u * u;
      dv1[i * ldv] = -(w * u) * inv_bn;
    }
    {
      const float u = t - v2[i * ldv];
      const float w = fabsf(tau - (u < 0.f ? 1.f : 0.f));
      l += w * u * u;
      dv2[i * ldv] = -(w * u) * inv_bn;
    }
  }
  const float bs = block_sum(l, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = bs;
  if (last_block_ticket(counter, &flag)) {
    float s = 0.f;
    for (int j = threadIdx.x; j < (int)gridDim.x; j += blockDim.x) s += __ldcg(partials + j);
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
      const float loss = s * inv_bn * 0.5f;
      loss_out[0] = loss;
      if (!isfinite(loss)) atomicOr(status, (uint32_t)PORL_FLAG_NONFINITE_LOSS);
    }
  }
}

// ---------------------------------------------------------------- K5
constexpr int kAwrWarps = 8;
constexpr int kAwrMaxD = 1024;

__global__ void __launch_bounds__(kAwrWarps * 32)
awr_nll_kernel(const float* __restrict__ mean, int64_t ld_mean, const float* __restrict__ log_std,
               const float* __restrict__ x, int64_t ld_x, const float* __restrict__ target_v,
               const float* __restrict__ v1n, const float* __restrict__ v2n, int64_t ldv, float alpha,
               int mode, int tanh_out, int64_t B, int D, float inv_bn, float* __restrict__ dz_out,
               float* __restrict__ dlog_std, float* __restrict__ loss_out, float* partials,
               uint32_t* counter, uint32_t* status) {
  pdl_entry();
  extern __shared__ float sm[];                 // [kAwrWarps][D] log_std-gradient accumulators
  __shared__ float red[32];
  __shared__ int flag;
  __shared__ float s_sum_ls;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* acc = sm + (size_t)warp * D;
  for (int j = lane; j < D; j += 32) acc[j] = 0.f;

  float ls_part = 0.f;
  for (int j = threadIdx.x; j < D; j += blockDim.x) ls_part += fminf(fmaxf(log_std[j], -5.f), 2.f);
  ls_part = block_sum(ls_part, red);
  if (threadIdx.x == 0) s_sum_ls = ls_part;
  __syncthreads();
  const float nll_const = s_sum_ls + 0.5f * (float)D * kLog2Pi;

  float loss_acc = 0.f;  // lane 0 of each warp
  for (int64_t i = (int64_t)blockIdx.x * kAwrWarps + warp; i < B; i += (int64_t)gridDim.x * kAwrWarps) {
    const float v = fminf(v1n[i * ldv], v2n[i * ldv]);
    const float adv = target_v[i] - v;
    const float e = (mode == PORL_AWR_POR) ? adv / alpha : alpha * adv;
    const float w = fminf(expf(e), kWeightCap);
    const float wn = w * inv_bn;
    const float* mrow = mean + i * ld_mean;
    const float* xrow = x + i * ld_x;
    float* drow = dz_out + i * ld_mean;
    float sz2 = 0.f;
    for (int j = lane; j < D; j += 32) {
      const float ls = fminf(fmaxf(log_std[j], -5.f), 2.f);
      const float inv_std = expf(-ls);
      const float mu = mrow[j];
      const float z = (xrow[j] - mu) * inv_std;
      sz2 = fmaf(z, z, sz2);
      float dmu = -wn * z * inv_std;
      if (tanh_out) dmu *= (1.f - mu * mu);
      drow[j] = dmu;
      acc[j] += wn * (1.f - z * z);
    }
    for (int j = D + lane; j < ld_mean; j += 32) drow[j] = 0.f;
    sz2 = warp_sum(sz2);
    const float nll = 0.5f * sz2 + nll_const;
    if (lane == 0) {
      if (!(nll > 0.f)) atomicOr(status, (uint32_t)PORL_FLAG_NONPOSITIVE_NLL);
      loss_acc += w * nll;
    }
  }
  __syncthreads();
  // block partials: column sums over the warps (fixed order), then the scalar loss
  float* bp = partials + (size_t)blockIdx.x * (D + 1);
  for (int j = threadIdx.x; j < D; j += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int wdx = 0; wdx < kAwrWarps; ++wdx) s += sm[(size_t)wdx * D + j];
    bp[j] = s;
  }
  const float bl = block_sum(lane == 0 ? loss_acc : 0.f, red);
  if (threadIdx.x == 0) bp[D] = bl;

  if (last_block_ticket(counter, &flag)) {
    // one warp per column: lanes take the block partials b = lane, lane + 32, ... in order, then a fixed shuffle
    // tree -- order-fixed like a serial sum, without its chain of gridDim.x dependent L2 loads
    const int nb = gridDim.x;
    for (int j = warp; j <= D; j += kAwrWarps) {
      float s = 0.f;
      for (int b = lane; b < nb; b += 32) s += __ldcg(partials + (size_t)b * (D + 1) + j);
      s = warp_sum(s);
      if (lane != 0) continue;
      if (j < D) {
        const float raw = log_std[j];
        dlog_std[j] = (raw >= -5.f && raw <= 2.f) ? s : 0.f;
      } else {
        const float loss = s * inv_bn;
        loss_out[0] = loss;
        if (!isfinite(loss)) atomicOr(status, (uint32_t)PORL_FLAG_NONFINITE_LOSS);
      }
    }
  }
}

// ---------------------------------------------------------------- K6
constexpr int kCqlMaxA = 32;

__global__ void __launch_bounds__(256)
cql_head_kernel(const float* __restrict__ q, const float* __restrict__ qn, int64_t ldq,
                const int64_t* __restrict__ actions, const float* __restrict__ rewards,
                const float* __restrict__ dones, float gamma, float alpha, int64_t B, int A,
                float inv_bn, float* __restrict__ dq, float* __restrict__ losses_out, float* partials,
                uint32_t* counter, uint32_t* status) {
  pdl_entry();
  __shared__ float red[32];
  __shared__ int flag;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float td_i = 0.f, pen_i = 0.f;
  if (i < B) {
    const float* qr = q + i * ldq;
    const float* nr = qn + i * ldq;
    float nmax = -INFINITY, qmax = -INFINITY;
    for (int a = 0; a < A; ++a) {
      nmax = fmaxf(nmax, nr[a]);
      qmax = fmaxf(qmax, qr[a]);
    }
    const float t = rewards[i] + (gamma * nmax) * (1.f - dones[i]);
    int64_t ai = actions[i];
    if (ai < 0 || ai >= A) {
      atomicOr(status, (uint32_t)PORL_FLAG_BAD_ACTION);
      ai = ai < 0 ? 0 : A - 1;
    }
    float se = 0.f;
    for (int a = 0; a < A; ++a) se += expf(qr[a] - qmax);
    const float lse = qmax + logf(se);
    const float qa = qr[ai];
    const float diff = qa - t;
    td_i = diff * diff;
    pen_i = lse - logf((float)A) - qa;
    float* dr = dq + i * ldq;
    for (int a = 0; a < A; ++a) {
      const float onehot = (a == (int)ai) ? 1.f : 0.f;
      dr[a] = inv_bn * (2.f * diff * onehot + alpha * (expf(qr[a] - lse) - onehot));
    }
    for (int a = A; a < ldq; ++a) dr[a] = 0.f;
  }
  const float bt = block_sum(td_i, red);
  const float bp = block_sum(pen_i, red);
  if (threadIdx.x == 0) {
    partials[2 * blockIdx.x] = bt;
    partials[2 * blockIdx.x + 1] = bp;
  }
  if (last_block_ticket(counter, &flag)) {
    float st = 0.f, sp = 0.f;
    for (int j = threadIdx.x; j < (int)gridDim.x; j += blockDim.x) {
      st += __ldcg(partials + 2 * j);
      sp += __ldcg(partials + 2 * j + 1);
    }
    st = block_sum(st, red);
    sp = block_sum(sp, red);
    if (threadIdx.x == 0) {
      const float td = st * inv_bn, pen = sp * inv_bn;
      const float loss = td + alpha * pen;
      losses_out[0] = loss;
      losses_out[1] = td;
      losses_out[2] = pen;
      if (!isfinite(loss)) atomicOr(status, (uint32_t)PORL_FLAG_NONFINITE_LOSS);
    }
  }
}

}  // namespace porl

using namespace porl;

extern "C" int porl_expectile_twin_loss(porl_ctx_t* ctx, const float* v1, const float* v2,
                                        const float* nv1, const float* nv2, int64_t ldv,
                                        const float* rewards, const float* dones, float gamma, float tau,
                                        int64_t B, int64_t B_norm, float* target_v, float* dv1,
                                        float* dv2, float* loss_out, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && v1 && v2 && nv1 && nv2 && rewards && dones && target_v && dv1 && dv2 && loss_out,
                 "porl_expectile_twin_loss: null argument");
  PORL_CHECK_ARG(B > 0 && B_norm >= B && ldv >= 1, "porl_expectile_twin_loss: bad B=%lld B_norm=%lld",
                 (long long)B, (long long)B_norm);
  const int64_t blocks = (B + 255) / 256;
  PORL_CHECK_ARG(blocks <= kPartialFloats, "porl_expectile_twin_loss: batch too large");
  ProfScope prof(ctx, PORL_CAT_LOSS, (double)B * 9.0 * 4.0, as_stream(stream));
  launch_pdl(ctx, expectile_kernel, (int)blocks, 256, 0, as_stream(stream), 
      v1, v2, nv1, nv2, ldv, rewards, dones, gamma, tau, B, 1.0f / (float)B_norm, target_v, dv1, dv2,
      loss_out, ctx->partials, ctx->counters + 1, ctx->counters);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

extern "C" int porl_awr_gauss_nll(porl_ctx_t* ctx, const float* mean, int64_t ld_mean,
                                  const float* log_std, const float* x, int64_t ld_x,
                                  const float* target_v, const float* v1n, const float* v2n, int64_t ldv,
                                  float alpha, int32_t mode, int32_t tanh_out, int64_t B, int32_t D,
                                  int64_t B_norm, float* dz_out, float* dlog_std, float* loss_out,
                                  porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && mean && log_std && x && target_v && v1n && v2n && dz_out && dlog_std && loss_out,
                 "porl_awr_gauss_nll: null argument");
  PORL_CHECK_ARG(B > 0 && B_norm >= B && D > 0 && D <= kAwrMaxD && ld_mean >= D && ld_x >= D && ldv >= 1 &&
                     (mode == PORL_AWR_POR || mode == PORL_AWR_SORL) && alpha != 0.f,
                 "porl_awr_gauss_nll: bad shape B=%lld D=%d ld_mean=%lld ld_x=%lld mode=%d",
                 (long long)B, D, (long long)ld_mean, (long long)ld_x, mode);
  int64_t blocks = (B + kAwrWarps - 1) / kAwrWarps;
  const int64_t cap_sm = (int64_t)ctx->sm_count * (D > 64 ? 4 : 1);   // narrow rows: fewer partials to sum at the end
  const int64_t cap_ws = kPartialFloats / (D + 1);
  if (blocks > cap_sm) blocks = cap_sm;
  if (blocks > cap_ws) blocks = cap_ws;
  const size_t smem = (size_t)kAwrWarps * D * sizeof(float);
  ProfScope prof(ctx, PORL_CAT_LOSS, (double)B * (3.0 * D + 3.0) * 4.0, as_stream(stream));
  launch_pdl(ctx, awr_nll_kernel, (int)blocks, kAwrWarps * 32, smem, as_stream(stream), 
      mean, ld_mean, log_std, x, ld_x, target_v, v1n, v2n, ldv, alpha, mode, tanh_out, B, D,
      1.0f / (float)B_norm, dz_out, dlog_std, loss_out, ctx->partials, ctx->counters + 2, ctx->counters);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

extern "C" int porl_cql_head(porl_ctx_t* ctx, const float* q, const float* q_next_tgt, int64_t ldq,
                             const int64_t* actions, const float* rewards, const float* dones,
                             float gamma, float alpha, int64_t B, int32_t A, int64_t B_norm, float* dq,
                             float* losses_out, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && q && q_next_tgt && actions && rewards && dones && dq && losses_out,
                 "porl_cql_head: null argument");
  PORL_CHECK_ARG(B > 0 && B_norm >= B && A > 0 && A <= kCqlMaxA && ldq >= A,
                 "porl_cql_head: bad shape B=%lld A=%d ldq=%lld", (long long)B, A, (long long)ldq);
  const int64_t blocks = (B + 255) / 256;
  PORL_CHECK_ARG(2 * blocks <= kPartialFloats, "porl_cql_head: batch too large");
  ProfScope prof(ctx, PORL_CAT_LOSS, (double)B * (3.0 * A * 4.0 + 16.0), as_stream(stream));
  launch_pdl(ctx, cql_head_kernel, (int)blocks, 256, 0, as_stream(stream), 
      q, q_next_tgt, ldq, actions, rewards, dones, gamma, alpha, B, A, 1.0f / (float)B_norm, dq,
      losses_out, ctx->partials, ctx->counters + 3, ctx->counters);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

// ---------------------------------------------------------------- no-allocation variant of K4
// Same kernel as porl_expectile_twin_loss, but the reduction scratch (ticket word, status word, block partials) lives
// in a file-scope device array instead of the context's cudaMalloc'ed buffers, and no context is needed: the entry
// point calls no cudaMalloc / cudaFree / allocating routine, so it can be used where allocation is forbidden (CUDA
// graph capture, serving loops).  Every GPU that loads the library holds the array (kNoAllocScratchBytes) until exit.
// All calls on a device SHARE the array: calls that overlap on the GPU (two streams) race.
namespace {

constexpr int kNoAllocMaxBlocks = PORL_NOALLOC_MAX_ROWS / 256;            // one partial per 256-row block
constexpr int kNoAllocHeadWords = 64;                                     // [0] ticket, [1] status, rest padding
constexpr size_t kNoAllocScratchBytes = (size_t)(kNoAllocHeadWords + kNoAllocMaxBlocks) * 4;
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) uint32_t g_noalloc_scratch[kNoAllocHeadWords + kNoAllocMaxBlocks];

// device address of the array per device ordinal, resolved once with that device current (the symbol's name is a
// host-side stand-in in host code, never a device address)
uint32_t* g_noalloc_ptr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mu;

int noalloc_scratch(uint32_t** out) {
  int dev = -1;
  PORL_CUDA(cudaGetDevice(&dev));
  PORL_CHECK_ARG(dev >= 0 && dev < kNoAllocMaxDevices, "porl_expectile_twin_loss_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    uint32_t* p = nullptr;
    PORL_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    PORL_CHECK_ARG(p != nullptr, "porl_expectile_twin_loss_noalloc: cudaGetSymbolAddress returned null");
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return PORL_OK;
}

}  // namespace

extern "C" int porl_expectile_twin_loss_noalloc(const float* v1, const float* v2, const float* nv1, const float* nv2,
                                                int64_t ldv, const float* rewards, const float* dones, float gamma,
                                                float tau, int64_t B, int64_t B_norm, float* target_v, float* dv1,
                                                float* dv2, float* loss_out, porl_stream_t stream) {
  PORL_CHECK_ARG(v1 && v2 && nv1 && nv2 && rewards && dones && target_v && dv1 && dv2 && loss_out,
                 "porl_expectile_twin_loss_noalloc: null argument");
  PORL_CHECK_ARG(B > 0 && B_norm >= B && ldv >= 1, "porl_expectile_twin_loss_noalloc: bad B=%lld B_norm=%lld",
                 (long long)B, (long long)B_norm);
  PORL_CHECK_ARG(B <= (int64_t)PORL_NOALLOC_MAX_ROWS,
                 "porl_expectile_twin_loss_noalloc: B=%lld exceeds the static scratch's maximum of %d rows "
                 "(PORL_NOALLOC_MAX_ROWS); use porl_expectile_twin_loss",
                 (long long)B, (int)PORL_NOALLOC_MAX_ROWS);
  uint32_t* scratch = nullptr;
  int rc = noalloc_scratch(&scratch);
  if (rc != PORL_OK) return rc;
  const int blocks = (int)((B + 255) / 256);
  // the array keeps the last call's contents: the ticket and status words must start at zero (an aborted launch
  // would leave the ticket mid-count); the partials are fully rewritten by the blocks of this launch
  PORL_CUDA(cudaMemsetAsync(scratch, 0, kNoAllocHeadWords * sizeof(uint32_t), as_stream(stream)));
  expectile_kernel<<<blocks, 256, 0, as_stream(stream)>>>(
      v1, v2, nv1, nv2, ldv, rewards, dones, gamma, tau, B, 1.0f / (float)B_norm, target_v, dv1, dv2, loss_out,
      reinterpret_cast<float*>(scratch + kNoAllocHeadWords), scratch, scratch + 1);
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    porl::set_error("porl_expectile_twin_loss_noalloc: kernel launch failed: %s", cudaGetErrorString(e));
    return PORL_ERR_CUDA;
  }
  return PORL_OK;
}
