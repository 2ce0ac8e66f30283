// K1: minibatch gather from the HBM-resident dataset (bit-exact payload copy).
//
// HBM-bound: one warp walks one (or more) rows; a row is 2.9 KB contiguous, so each warp load
// instruction is a fully used 128 B line (512 B when the row base allows float4).  The AoS variant
// fuses the column split of por_train.py:74-78 and writes the padded SoA minibatch directly, so
// the batch is written once and never re-sliced.
#include "common.cuh"

namespace porl {

__device__ __forceinline__ int64_t checked_index(const int64_t* idx, int64_t b, int64_t n_rows,
                                                 uint32_t* status) {
  int64_t r = idx[b];
  if (r < 0 || r >= n_rows) {
    atomicOr(status, (uint32_t)PORL_FLAG_BAD_INDEX);
    r = r < 0 ? 0 : n_rows - 1;
  }
  return r;
}

// copy n floats src -> dst with one warp; float4 when both are 16-byte aligned
__device__ __forceinline__ void warp_copy(const float* __restrict__ src, float* __restrict__ dst,
                                          int n, int lane) {
  const bool v4 = (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15u) == 0);
  if (v4) {
    const int n4 = n >> 2;
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* d4 = reinterpret_cast<float4*>(dst);
    for (int i = lane; i < n4; i += 32) d4[i] = __ldg(s4 + i);
    for (int i = (n4 << 2) + lane; i < n; i += 32) dst[i] = __ldg(src + i);
  } else {
    int i = lane;
    for (; i + 96 < n; i += 128) {
      const float a = __ldg(src + i), b = __ldg(src + i + 32), c = __ldg(src + i + 64), d = __ldg(src + i + 96);
      dst[i] = a; dst[i + 32] = b; dst[i + 64] = c; dst[i + 96] = d;
    }
    for (; i < n; i += 32) dst[i] = __ldg(src + i);
  }
}

__global__ void __launch_bounds__(256)
gather_rows_kernel(const float* __restrict__ table, int64_t n_rows, int64_t row_stride, int row_len,
                   const int64_t* __restrict__ idx, int64_t B, float* __restrict__ out,
                   int64_t out_stride, uint32_t* status) {
  pdl_entry();
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t b = warp; b < B; b += nwarps) {
    const int64_t r = checked_index(idx, b, n_rows, status);
    warp_copy(table + r * row_stride, out + b * out_stride, row_len, lane);
  }
}

__global__ void __launch_bounds__(256)
gather_transitions_kernel(const float* __restrict__ table, int64_t n_rows, int64_t row_stride, int S,
                          int A, const int64_t* __restrict__ idx, int64_t B, float* __restrict__ obs,
                          float* __restrict__ next_obs, int64_t ld_obs, float* __restrict__ rewards,
                          float* __restrict__ dones, float* __restrict__ actions, int64_t ld_act,
                          uint32_t* status) {
  pdl_entry();
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t b = warp; b < B; b += nwarps) {
    const int64_t r = checked_index(idx, b, n_rows, status);
    const float* row = table + r * row_stride;
    float* o = obs + b * ld_obs;
    float* n = next_obs + b * ld_obs;
    warp_copy(row, o, S, lane);
    warp_copy(row + S + 1, n, S, lane);
    for (int i = S + lane; i < ld_obs; i += 32) {
      o[i] = 0.f;
      n[i] = 0.f;
    }
    if (lane == 0) {
      rewards[b] = __ldg(row + S);
      dones[b] = __ldg(row + 2 * S + 1);
    }
    if (lane < A) actions[b * ld_act + lane] = __ldg(row + 2 * S + 2 + lane);
    for (int i = A + lane; i < ld_act; i += 32) actions[b * ld_act + i] = 0.f;
  }
}

__global__ void __launch_bounds__(256)
gather_soa_kernel(const float* __restrict__ states, const float* __restrict__ next_states,
                  const int64_t* __restrict__ actions, const float* __restrict__ rewards,
                  const float* __restrict__ dones, int64_t cap, int S, int64_t ld_state,
                  const int64_t* __restrict__ idx, int64_t B, float* __restrict__ o_states,
                  float* __restrict__ o_next, int64_t ld_out, int64_t* __restrict__ o_actions,
                  float* __restrict__ o_rewards, float* __restrict__ o_dones, uint32_t* status) {
  pdl_entry();
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t b = warp; b < B; b += nwarps) {
    const int64_t r = checked_index(idx, b, cap, status);
    float* o = o_states + b * ld_out;
    float* n = o_next + b * ld_out;
    warp_copy(states + r * ld_state, o, S, lane);
    warp_copy(next_states + r * ld_state, n, S, lane);
    for (int i = S + lane; i < ld_out; i += 32) {
      o[i] = 0.f;
      n[i] = 0.f;
    }
    if (lane == 0) {
      o_actions[b] = actions[r];
      o_rewards[b] = rewards[r];
      o_dones[b] = dones[r];
    }
  }
}

__global__ void __launch_bounds__(256)
pack_rows_kernel(const float* __restrict__ src, int64_t ld_src, int64_t B, int n,
                 float* __restrict__ dst, int64_t ld_dst) {
  pdl_entry();
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t b = warp; b < B; b += nwarps) {
    float* d = dst + b * ld_dst;
    warp_copy(src + b * ld_src, d, n, lane);
    for (int i = n + lane; i < ld_dst; i += 32) d[i] = 0.f;
  }
}

static int gather_grid(const porl_ctx* ctx, int64_t B) {
  // 8 warps per CTA, one row per warp, capped at 16 CTAs per SM (grid-stride beyond that)
  int64_t blocks = (B + 7) / 8;
  const int64_t cap = (int64_t)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace porl

using namespace porl;

extern "C" int porl_gather_rows_f32(porl_ctx_t* ctx, const float* table, int64_t n_rows,
                                    int64_t row_stride, int64_t row_len, const int64_t* idx,
                                    int64_t B, float* out, int64_t out_stride, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && table && idx && out, "porl_gather_rows_f32: null argument");
  PORL_CHECK_ARG(n_rows > 0 && row_len > 0 && row_len <= row_stride && row_len <= out_stride && B >= 0 &&
                     row_len < (1 << 30),
                 "porl_gather_rows_f32: bad shape n_rows=%lld row_len=%lld row_stride=%lld out_stride=%lld",
                 (long long)n_rows, (long long)row_len, (long long)row_stride, (long long)out_stride);
  if (B == 0) return PORL_OK;
  ProfScope prof(ctx, PORL_CAT_GATHER, (double)B * (8.0 * row_len + 8.0), as_stream(stream));
  launch_pdl(ctx, gather_rows_kernel, gather_grid(ctx, B), 256, 0, as_stream(stream), 
      table, n_rows, row_stride, (int)row_len, idx, B, out, out_stride, ctx->counters);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

extern "C" int porl_gather_transitions(porl_ctx_t* ctx, const float* table, int64_t n_rows,
                                       int64_t row_stride, int32_t S, int32_t A, const int64_t* idx,
                                       int64_t B, float* obs, float* next_obs, int64_t ld_obs,
                                       float* rewards, float* dones, float* actions, int64_t ld_act,
                                       porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && table && idx && obs && next_obs && rewards && dones && actions,
                 "porl_gather_transitions: null argument");
  PORL_CHECK_ARG(S > 0 && A >= 0 && A <= 32 && row_stride >= 2 * (int64_t)S + 2 + A && ld_obs >= S &&
                     ld_act >= A && n_rows > 0 && B >= 0,
                 "porl_gather_transitions: bad shape S=%d A=%d row_stride=%lld ld_obs=%lld ld_act=%lld",
                 S, A, (long long)row_stride, (long long)ld_obs, (long long)ld_act);
  if (B == 0) return PORL_OK;
  ProfScope prof(ctx, PORL_CAT_GATHER, (double)B * (8.0 * (2.0 * S + 2 + A) + 8.0), as_stream(stream));
  launch_pdl(ctx, gather_transitions_kernel, gather_grid(ctx, B), 256, 0, as_stream(stream), 
      table, n_rows, row_stride, S, A, idx, B, obs, next_obs, ld_obs, rewards, dones, actions, ld_act,
      ctx->counters);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

extern "C" int porl_gather_soa(porl_ctx_t* ctx, const float* states, const float* next_states,
                               const int64_t* actions, const float* rewards, const float* dones,
                               int64_t cap, int32_t S, int64_t ld_state, const int64_t* idx, int64_t B,
                               float* out_states, float* out_next_states, int64_t ld_out,
                               int64_t* out_actions, float* out_rewards, float* out_dones,
                               porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && states && next_states && actions && rewards && dones && idx && out_states &&
                     out_next_states && out_actions && out_rewards && out_dones,
                 "porl_gather_soa: null argument");
  PORL_CHECK_ARG(S > 0 && ld_state >= S && ld_out >= S && cap > 0 && B >= 0,
                 "porl_gather_soa: bad shape S=%d ld_state=%lld ld_out=%lld cap=%lld", S,
                 (long long)ld_state, (long long)ld_out, (long long)cap);
  if (B == 0) return PORL_OK;
  ProfScope prof(ctx, PORL_CAT_GATHER, (double)B * (8.0 * (2.0 * S + 2) + 16.0 + 8.0), as_stream(stream));
  launch_pdl(ctx, gather_soa_kernel, gather_grid(ctx, B), 256, 0, as_stream(stream), 
      states, next_states, actions, rewards, dones, cap, S, ld_state, idx, B, out_states,
      out_next_states, ld_out, out_actions, out_rewards, out_dones, ctx->counters);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

extern "C" int porl_pack_rows(porl_ctx_t* ctx, const float* src, int64_t ld_src, int64_t B, int32_t n,
                              float* dst, int64_t ld_dst, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && src && dst, "porl_pack_rows: null argument");
  PORL_CHECK_ARG(n > 0 && ld_src >= n && ld_dst >= n && B >= 0, "porl_pack_rows: bad shape n=%d ld_src=%lld ld_dst=%lld",
                 n, (long long)ld_src, (long long)ld_dst);
  if (B == 0) return PORL_OK;
  ProfScope prof(ctx, PORL_CAT_GATHER, (double)B * 8.0 * n, as_stream(stream));
  launch_pdl(ctx, pack_rows_kernel, gather_grid(ctx, B), 256, 0, as_stream(stream), src, ld_src, B, n, dst, ld_dst);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}
