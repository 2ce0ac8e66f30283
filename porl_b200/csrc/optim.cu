// K7: fused Adam step over a flat parameter buffer, with the Polyak / hard target update folded in
// and the optimizer clock (step count, cosine-annealed LR) kept on the device.
//
// HBM-bound: 28 B/param (read p,g,m,v; write p,m,v) + 8 B/param when a target net rides along.
// float4 accesses, grid-stride, one pass.  The scalar maths follows torch/optim/adam.py
// (_single_tensor_adam): bias corrections and step size in double, per-element work in fp32.
#include <math.h>
#include <string.h>

#include "common.cuh"
#include "optim.cuh"

namespace porl {

// ---- peer-memory gradient sum (data parallel): signalling words live in every rank's pad ----
struct PeerArgs {
  int world, rank;
  const float* bucket[PORL_MAX_PEERS];
  uint32_t* pad[PORL_MAX_PEERS];
  uint32_t* epoch;
  float* tail_out;
  int tail_floats;
  unsigned long long wait_ns;
  uint32_t* status;
};

__device__ __forceinline__ void peer_signal(uint32_t* addr, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t peer_poll(const uint32_t* addr) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// spins until *addr has reached `epoch` (wrap-safe); gives up after wait_ns, raises the status flag and returns false
__device__ __forceinline__ bool peer_wait(const uint32_t* addr, uint32_t epoch, unsigned long long wait_ns,
                                          uint32_t* status) {
  if ((int)(peer_poll(addr) - epoch) >= 0) return true;
  const unsigned long long t0 = global_ns();
  while ((int)(peer_poll(addr) - epoch) < 0) {
    __nanosleep(64);
    if (global_ns() - t0 > wait_ns) {
      atomicOr(status, (uint32_t)PORL_FLAG_PEER_TIMEOUT);
      return false;
    }
  }
  return true;
}
// gradients of the peers change between launches at fixed addresses: read them around L1
__device__ __forceinline__ float4 peer_load4(const float* base, int64_t i4) {
  return __ldcv(reinterpret_cast<const float4*>(base) + i4);
}

template <bool PEER>
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
            float* __restrict__ v, int64_t n, porl_opt_state_t* st, float beta1, float beta2, float eps,
            float* __restrict__ target, float ema_beta, int hard_every, const PeerArgs peers) {
  pdl_entry();
  __shared__ AdamScalars sh;
  __shared__ uint32_t sh_epoch;
  __shared__ int sh_timeout;
  __shared__ long long sh_t;
  __shared__ double sh_lr;
  // the LAST block of the grid owns no element: it publishes the scalars of the next update (optim.cuh) while the
  // other blocks work, and takes its ticket like them
  const int n_work = (int)gridDim.x - 1;
  const bool publisher = (int)blockIdx.x == n_work;
  const bool ema = (target != nullptr) && ema_beta > 0.f;
  const int64_t n4 = n >> 2;
  const int64_t stride = (int64_t)n_work * blockDim.x;
  float4* p4 = reinterpret_cast<float4*>(p);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  float4* t4 = reinterpret_cast<float4*>(target);
  // the first element's loads are issued before the (double-precision) bias corrections of thread 0, so every
  // block's HBM latency overlaps that scalar work instead of following it
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float4 pp0 = make_float4(0.f, 0.f, 0.f, 0.f), mm0 = pp0, vv0 = pp0, gg0 = pp0, tt0 = pp0;
  if (!publisher && i0 < n4) {
    pp0 = p4[i0]; mm0 = m4[i0]; vv0 = v4[i0];
    if (!PEER) gg0 = g4[i0];
    if (ema) tt0 = t4[i0];
  }
  if (threadIdx.x == 0) {
    long long t_now;
    double lr_now;
    sh = adam_scalars_fast(st, beta1, beta2, hard_every, &t_now, &lr_now);
    sh_t = t_now;
    sh_lr = lr_now;
    if (PEER) sh_epoch = *reinterpret_cast<volatile uint32_t*>(peers.epoch) + 1u;
    sh_timeout = 0;
  }
  __syncthreads();
  if (publisher && threadIdx.x == 32) adam_publish_next(st, sh_t, sh_lr, beta1, beta2);
  const AdamScalars s = sh;
  bool skip = false;
  if (PEER) {
    // 1. "my bucket of this epoch is complete" -> every rank's pad (own one included); every block then waits for
    //    all ranks' words in ITS OWN rank's pad, so no block depends on another block of the same grid.
    const uint32_t e = sh_epoch;
    if (blockIdx.x == 0 && threadIdx.x < peers.world) {
      __threadfence_system();
      peer_signal(peers.pad[threadIdx.x] + peers.rank, e);
    }
    if (threadIdx.x < peers.world && !peer_wait(peers.pad[peers.rank] + threadIdx.x, e, peers.wait_ns, peers.status))
      sh_timeout = 1;
    __syncthreads();
    // A peer that never signalled: its bucket is not the gradient of this update.  The launch then leaves parameters,
    // moments, target net and the optimizer clock untouched (every block decides from the device status word, which
    // stays set until the host reads it, so blocks that started after the flag was raised skip as well) -- the step is
    // lost, not corrupted, and the next loss read raises PORL_FLAG_PEER_TIMEOUT.
    skip = sh_timeout || (*reinterpret_cast<volatile uint32_t*>(peers.status) & (uint32_t)PORL_FLAG_PEER_TIMEOUT);
    if (!skip && blockIdx.x == 0 && threadIdx.x < peers.tail_floats) {
      float t = 0.f;
      for (int r = 0; r < peers.world; ++r) t += __ldcv(peers.bucket[r] + n + threadIdx.x);
      peers.tail_out[threadIdx.x] = t;
    }
  }
  const bool hard = (target != nullptr) && s.do_hard;
  const float keep = 1.f - ema_beta;

  for (int64_t i = i0; i < n4 && !skip && !publisher; i += stride) {
    float4 pp, mm, vv;
    if (i == i0) { pp = pp0; mm = mm0; vv = vv0; }
    else { pp = p4[i]; mm = m4[i]; vv = v4[i]; }
    float4 gg;
    if (PEER) {
      float4 gr[PORL_MAX_PEERS];
#pragma unroll
      for (int r = 0; r < PORL_MAX_PEERS; ++r)
        if (r < peers.world) gr[r] = peer_load4(peers.bucket[r], i);
      gg = gr[0];
#pragma unroll
      for (int r = 1; r < PORL_MAX_PEERS; ++r)
        if (r < peers.world) { gg.x += gr[r].x; gg.y += gr[r].y; gg.z += gr[r].z; gg.w += gr[r].w; }
    } else {
      gg = (i == i0) ? gg0 : g4[i];
    }
    adam_elem(pp.x, gg.x, mm.x, vv.x, s, beta1, beta2, eps);
    adam_elem(pp.y, gg.y, mm.y, vv.y, s, beta1, beta2, eps);
    adam_elem(pp.z, gg.z, mm.z, vv.z, s, beta1, beta2, eps);
    adam_elem(pp.w, gg.w, mm.w, vv.w, s, beta1, beta2, eps);
    p4[i] = pp; m4[i] = mm; v4[i] = vv;
    if (ema) {
      float4 tt = (i == i0) ? tt0 : t4[i];
      tt.x = fmaf(ema_beta, pp.x, tt.x * keep);       // mul_(1-beta).add_(src, alpha=beta)
      tt.y = fmaf(ema_beta, pp.y, tt.y * keep);
      tt.z = fmaf(ema_beta, pp.z, tt.z * keep);
      tt.w = fmaf(ema_beta, pp.w, tt.w * keep);
      t4[i] = tt;
    } else if (hard) {
      t4[i] = pp;
    }
  }
  // tail (n not a multiple of 4)
  for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n && !skip && !publisher; i += stride) {
    float pp = p[i], mm = m[i], vv = v[i];
    float gi;
    if (PEER) {
      gi = __ldcv(peers.bucket[0] + i);
      for (int r = 1; r < peers.world; ++r) gi += __ldcv(peers.bucket[r] + i);
    } else {
      gi = g[i];
    }
    adam_elem(pp, gi, mm, vv, s, beta1, beta2, eps);
    p[i] = pp; m[i] = mm; v[i] = vv;
    if (ema) target[i] = fmaf(ema_beta, pp, target[i] * keep);
    else if (hard) target[i] = pp;
  }

  // the last block to finish advances the optimizer clock
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t ticket = atomicAdd(&st->blocks_done, 1u);
    if (ticket == gridDim.x - 1 && skip) st->blocks_done = 0;
    if (ticket == gridDim.x - 1 && !skip) {
      st->blocks_done = 0;
      adam_advance_clock(st);
      if (PEER) *reinterpret_cast<volatile uint32_t*>(peers.epoch) = sh_epoch;
      __threadfence();
    }
    if (PEER) sh.do_hard = (ticket == gridDim.x - 1) && !skip;      // reuse as "this block closes the launch"
  }
  if (PEER) {
    // 3. the closing block: "I have read your buckets" to every rank, then wait until every rank has read mine
    __syncthreads();
    if (sh.do_hard && threadIdx.x < peers.world) {
      const uint32_t e = sh_epoch;
      __threadfence_system();
      peer_signal(peers.pad[threadIdx.x] + PORL_MAX_PEERS + peers.rank, e);
      peer_wait(peers.pad[peers.rank] + PORL_MAX_PEERS + threadIdx.x, e, peers.wait_ns, peers.status);
    }
  }
}


// ---- two-shot (sharded) data-parallel optimizer step ------------------------------------------------------------
// Rank r owns shard r of the flat group: it sums that shard of the gradient over all ranks' buckets (rank order), runs
// Adam / Polyak on it and STORES the new parameters (and target-net values) into every rank's buffers over NVLink.
// Per rank and step: (world-1)/world * n floats read + (world-1)/world * n (x2 with a target net) written over the
// link instead of (world-1) * n read, and 1/world of the Adam HBM traffic.  Every element is computed by exactly one
// rank and broadcast, so the replicas are bitwise identical by construction.  The closing handshake ("I have read your
// bucket and written my shard into your buffers") keeps the kernel alive until every peer's stores have landed here.
struct ShardRank {
  float* p;                         // this rank's parameters (== params[rank])
  float* m;
  float* v;
  float* target;                    // this rank's target net or null (== targets[rank])
  porl_opt_state_t* st;
  float* params[PORL_MAX_PEERS];    // every rank's parameter buffer as mapped by this rank
  float* targets[PORL_MAX_PEERS];
  PeerArgs peers;
};

__global__ void __launch_bounds__(256)
adam_sharded_kernel(const __grid_constant__ ShardRank self, const ShardRank* __restrict__ emu, int64_t n, float beta1,
                    float beta2, float eps, float ema_beta, int hard_every) {
  pdl_entry();
  // blockIdx.y selects the rank when several ranks are emulated inside one cooperative launch (single-GPU tests)
  __shared__ ShardRank a;
  __shared__ AdamScalars sh;
  __shared__ long long sh_t;
  __shared__ double sh_lr;
  __shared__ uint32_t sh_epoch;
  __shared__ int sh_flag;
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(emu ? emu + blockIdx.y : &self);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&a);
    for (int i = threadIdx.x; i < (int)(sizeof(ShardRank) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const int world = a.peers.world, rank = a.peers.rank;
  if (threadIdx.x == 0) {
    long long t_now;
    double lr_now;
    sh = adam_scalars_fast(a.st, beta1, beta2, hard_every, &t_now, &lr_now);
    sh_t = t_now;
    sh_lr = lr_now;
    sh_epoch = *reinterpret_cast<volatile uint32_t*>(a.peers.epoch) + 1u;
    sh_flag = 0;
  }
  __syncthreads();
  const AdamScalars s = sh;
  const uint32_t e = sh_epoch;
  // 1. "my bucket of this epoch is complete" to every rank; every block waits for all ranks' words in its own pad
  if (blockIdx.x == 0 && threadIdx.x < world) {
    __threadfence_system();
    peer_signal(a.peers.pad[threadIdx.x] + rank, e);
  }
  if (threadIdx.x < world && !peer_wait(a.peers.pad[rank] + threadIdx.x, e, a.peers.wait_ns, a.peers.status)) sh_flag = 1;
  __syncthreads();
  const bool skip = sh_flag || (*reinterpret_cast<volatile uint32_t*>(a.peers.status) & (uint32_t)PORL_FLAG_PEER_TIMEOUT);
  if (!skip && blockIdx.x == 0 && threadIdx.x < a.peers.tail_floats) {
    float t = 0.f;
    for (int r = 0; r < world; ++r) t += __ldcv(a.peers.bucket[r] + n + threadIdx.x);
    a.peers.tail_out[threadIdx.x] = t;
  }
  // 2. this rank's shard: float4 range [s0, s1), chunks of a multiple of 32 float4 (512-byte aligned)
  const int64_t n4 = n >> 2;
  const int64_t chunk = ((n4 + world - 1) / world + 31) / 32 * 32;
  const int64_t s0 = (int64_t)rank * chunk < n4 ? (int64_t)rank * chunk : n4;
  const int64_t s1 = s0 + chunk < n4 ? s0 + chunk : n4;
  const bool ema = (a.target != nullptr) && ema_beta > 0.f;
  const bool hard = (a.target != nullptr) && s.do_hard;
  const float keep = 1.f - ema_beta;
  float4* m4 = reinterpret_cast<float4*>(a.m);
  float4* v4 = reinterpret_cast<float4*>(a.v);
  const float4* p4 = reinterpret_cast<const float4*>(a.p);
  const float4* t4 = reinterpret_cast<const float4*>(a.target);
  // the last block (in x) owns no element: it publishes the scalars of the next update while the others work (optim.cuh)
  const int n_work = (int)gridDim.x - 1;
  const bool publisher = (int)blockIdx.x == n_work;
  if (publisher && threadIdx.x == 32) adam_publish_next(a.st, sh_t, sh_lr, beta1, beta2);
  const int64_t stride = (int64_t)n_work * blockDim.x;
  for (int64_t i = s0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < s1 && !skip && !publisher; i += stride) {
    float4 pp = p4[i], mm = m4[i], vv = v4[i];
    float4 gr[PORL_MAX_PEERS];
#pragma unroll
    for (int r = 0; r < PORL_MAX_PEERS; ++r)
      if (r < world) gr[r] = peer_load4(a.peers.bucket[r], i);
    float4 gg = gr[0];
#pragma unroll
    for (int r = 1; r < PORL_MAX_PEERS; ++r)
      if (r < world) { gg.x += gr[r].x; gg.y += gr[r].y; gg.z += gr[r].z; gg.w += gr[r].w; }
    adam_elem(pp.x, gg.x, mm.x, vv.x, s, beta1, beta2, eps);
    adam_elem(pp.y, gg.y, mm.y, vv.y, s, beta1, beta2, eps);
    adam_elem(pp.z, gg.z, mm.z, vv.z, s, beta1, beta2, eps);
    adam_elem(pp.w, gg.w, mm.w, vv.w, s, beta1, beta2, eps);
    m4[i] = mm; v4[i] = vv;
    float4 tt = pp;
    if (ema) {
      tt = t4[i];
      tt.x = fmaf(ema_beta, pp.x, tt.x * keep);
      tt.y = fmaf(ema_beta, pp.y, tt.y * keep);
      tt.z = fmaf(ema_beta, pp.z, tt.z * keep);
      tt.w = fmaf(ema_beta, pp.w, tt.w * keep);
    }
#pragma unroll
    for (int r = 0; r < PORL_MAX_PEERS; ++r) {
      if (r < world) {
        reinterpret_cast<float4*>(a.params[r])[i] = pp;
        if (ema || hard) reinterpret_cast<float4*>(a.targets[r])[i] = tt;
      }
    }
  }
  // 3. peer stores of this block visible system-wide, then the ticket; the last block advances the clock and closes
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    porl_opt_state_t* st = a.st;
    const uint32_t ticket = atomicAdd(&st->blocks_done, 1u);
    const bool last = ticket == gridDim.x - 1;
    if (last) {
      st->blocks_done = 0;
      if (!skip) {
        const long long ep = st->step + 1;
        if (st->t_max > 0) {
          const double T = (double)st->t_max;
          double lr = st->lr;
          if ((ep - 1 - st->t_max) % (2 * st->t_max) == 0) lr = lr + st->base_lr * (1.0 - cos(M_PI / T)) / 2.0;
          else lr = (1.0 + cos(M_PI * (double)ep / T)) / (1.0 + cos(M_PI * (double)(ep - 1) / T)) * lr;
          st->lr = lr;
        }
        st->step = ep;
        *reinterpret_cast<volatile uint32_t*>(a.peers.epoch) = e;
        __threadfence();
      }
    }
    sh_flag = (last && !skip) ? 2 : 0;
  }
  __syncthreads();
  if (sh_flag == 2 && threadIdx.x < world) {
    __threadfence_system();
    peer_signal(a.peers.pad[threadIdx.x] + PORL_MAX_PEERS + rank, e);
    peer_wait(a.peers.pad[rank] + PORL_MAX_PEERS + threadIdx.x, e, a.peers.wait_ns, a.peers.status);
  }
}

}  // namespace porl

using namespace porl;

extern "C" int porl_adam_step(porl_ctx_t* ctx, float* p, const float* g, float* m, float* v, int64_t n,
                              porl_opt_state_t* state_dev, float beta1, float beta2, float eps,
                              float* target, float ema_beta, int32_t hard_every, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && p && g && m && v && state_dev, "porl_adam_step: null argument");
  PORL_CHECK_ARG(n > 0, "porl_adam_step: n=%lld", (long long)n);
  PORL_CHECK_ARG(aligned16(p) && aligned16(g) && aligned16(m) && aligned16(v) && (!target || aligned16(target)),
                 "porl_adam_step: buffers must be 16-byte aligned");
  PORL_CHECK_ARG(!(ema_beta > 0.f && hard_every > 0), "porl_adam_step: choose Polyak or hard target update");
  int64_t blocks = ((n >> 2) + 255) / 256;
  // one wave: every block pays the clock read, a fence and a ticket, so a second wave of blocks costs more than a second
  // trip of the grid-stride loop (the kernel holds 3 blocks of 72 registers per SM; measured with ncu: 2.03 waves at 901
  // blocks, `barrier` and `membar` the top stalls)
  static int per_sm = 0;
  if (per_sm == 0) {
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, adam_kernel<false>, 256, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
  }
  const int64_t cap = (int64_t)ctx->sm_count * per_sm - 1;      // - the publisher block
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  ProfScope prof(ctx, PORL_CAT_OPTIMIZER, (double)n * (target ? 36.0 : 28.0), as_stream(stream));
  launch_pdl(ctx, adam_kernel<false>, (int)blocks + 1 /* publisher block */, 256, 0, as_stream(stream), p, g, m, v, n, state_dev, beta1, beta2, eps,
                                                                 target, ema_beta, hard_every, PeerArgs{});
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}


static int fill_shard_rank(porl_ctx_t* ctx, ShardRank& r, float* p, const porl_peer_group_t* pg, float* m, float* v,
                           porl_opt_state_t* st, float* target, int64_t n) {
  PORL_CHECK_ARG(p && pg && m && v && st, "porl_adam_step_peer: null argument");
  PORL_CHECK_ARG(n > 0 && (n & 3) == 0, "porl_adam_step_peer: n=%lld must be a positive multiple of 4", (long long)n);
  PORL_CHECK_ARG(pg->world >= 1 && pg->world <= PORL_MAX_PEERS && pg->rank >= 0 && pg->rank < pg->world,
                 "porl_adam_step_peer: world=%d rank=%d", pg->world, pg->rank);
  PORL_CHECK_ARG(pg->epoch && pg->tail_floats >= 0 && pg->tail_floats <= 256 && (pg->tail_floats == 0 || pg->tail_out),
                 "porl_adam_step_peer: epoch / tail_out missing or tail_floats=%d outside [0, 256]", pg->tail_floats);
  PORL_CHECK_ARG(aligned16(p) && aligned16(m) && aligned16(v) && (!target || aligned16(target)),
                 "porl_adam_step_peer: buffers must be 16-byte aligned");
  r = ShardRank{};
  r.p = p; r.m = m; r.v = v; r.target = target; r.st = st;
  PeerArgs& a = r.peers;
  a.world = pg->world;
  a.rank = pg->rank;
  for (int k = 0; k < pg->world; ++k) {
    PORL_CHECK_ARG(pg->bucket[k] && pg->pad[k] && aligned16(pg->bucket[k]), "porl_adam_step_peer: bucket/pad of rank %d", k);
    PORL_CHECK_ARG(pg->params[k] && aligned16(pg->params[k]) && (!target || (pg->target[k] && aligned16(pg->target[k]))),
                   "porl_adam_step_peer (sharded): parameter / target mapping of rank %d missing", k);
    a.bucket[k] = pg->bucket[k];
    a.pad[k] = pg->pad[k];
    r.params[k] = pg->params[k];
    r.targets[k] = pg->target[k];
  }
  PORL_CHECK_ARG(pg->params[pg->rank] == p && (!target || pg->target[pg->rank] == target),
                 "porl_adam_step_peer (sharded): params[rank] / target[rank] must be this rank's own buffers");
  a.epoch = pg->epoch;
  a.tail_out = pg->tail_out;
  a.tail_floats = pg->tail_floats;
  a.wait_ns = (unsigned long long)(pg->wait_ms > 0 ? pg->wait_ms : 20000) * 1000000ull;
  a.status = ctx->counters;
  return PORL_OK;
}

static int sharded_blocks(const porl_ctx_t* ctx, int64_t n, int world) {
  const int64_t n4 = n >> 2;
  const int64_t chunk = ((n4 + world - 1) / world + 31) / 32 * 32;
  int64_t blocks = (chunk + 255) / 256;
  const int64_t cap = (int64_t)ctx->sm_count * 4;
  if (blocks > cap) blocks = cap;
  return (int)(blocks < 1 ? 1 : blocks) + 1;      // + the publisher block (owns no element)
}

// Single-GPU stand-in for `world` ranks (tests): ONE cooperative launch whose blockIdx.y is the rank, so the blocks that
// wait on one another are co-resident by construction (separate launches that spin on each other are not).
extern "C" int porl_adam_step_peer_emulated(porl_ctx_t* ctx, int32_t world, const porl_peer_rank_t* ranks, int64_t n,
                                            float beta1, float beta2, float eps, float ema_beta, int32_t hard_every,
                                            porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && ranks && world >= 1 && world <= PORL_MAX_PEERS, "porl_adam_step_peer_emulated: world=%d", world);
  PORL_CHECK_ARG(!(ema_beta > 0.f && hard_every > 0), "porl_adam_step_peer_emulated: choose Polyak or hard target update");
  ShardRank host[PORL_MAX_PEERS];
  for (int r = 0; r < world; ++r) {
    PORL_CHECK_ARG(ranks[r].peers && ranks[r].peers->world == world && ranks[r].peers->rank == r,
                   "porl_adam_step_peer_emulated: rank %d's peer group", r);
    int rc = fill_shard_rank(ctx, host[r], ranks[r].p, ranks[r].peers, ranks[r].m, ranks[r].v, ranks[r].state,
                             ranks[r].target, n);
    if (rc != PORL_OK) return rc;
  }
  static thread_local ShardRank* dev = nullptr;
  if (!dev) PORL_CUDA(cudaMalloc(&dev, sizeof(ShardRank) * PORL_MAX_PEERS));
  cudaStream_t st = as_stream(stream);
  PORL_CUDA(cudaStreamSynchronize(st));
  PORL_CUDA(cudaMemcpy(dev, host, sizeof(ShardRank) * world, cudaMemcpyHostToDevice));
  const int blocks = sharded_blocks(ctx, n, world);
  int per_sm = 0;
  PORL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, adam_sharded_kernel, 256, 0));
  PORL_CHECK_ARG((int64_t)blocks * world <= (int64_t)per_sm * ctx->sm_count,
                 "porl_adam_step_peer_emulated: %d x %d blocks cannot be co-resident", blocks, world);
  ShardRank self = host[0];
  const ShardRank* emu = dev;
  void* args[] = {&self, &emu, &n, &beta1, &beta2, &eps, &ema_beta, &hard_every};
  PORL_CUDA(cudaLaunchCooperativeKernel((const void*)adam_sharded_kernel, dim3(blocks, world), dim3(256), args, 0, st));
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

extern "C" int porl_adam_step_peer(porl_ctx_t* ctx, float* p, const porl_peer_group_t* pg, float* m, float* v,
                                   int64_t n, porl_opt_state_t* state_dev, float beta1, float beta2, float eps,
                                   float* target, float ema_beta, int32_t hard_every, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  PORL_CHECK_ARG(ctx && p && pg && m && v && state_dev, "porl_adam_step_peer: null argument");
  PORL_CHECK_ARG(n > 0 && (n & 3) == 0, "porl_adam_step_peer: n=%lld must be a positive multiple of 4", (long long)n);
  PORL_CHECK_ARG(pg->world >= 1 && pg->world <= PORL_MAX_PEERS && pg->rank >= 0 && pg->rank < pg->world,
                 "porl_adam_step_peer: world=%d rank=%d", pg->world, pg->rank);
  PORL_CHECK_ARG(pg->epoch && pg->tail_floats >= 0 && pg->tail_floats <= 256 && (pg->tail_floats == 0 || pg->tail_out),
                 "porl_adam_step_peer: epoch / tail_out missing or tail_floats=%d outside [0, 256]", pg->tail_floats);
  PORL_CHECK_ARG(aligned16(p) && aligned16(m) && aligned16(v) && (!target || aligned16(target)),
                 "porl_adam_step_peer: buffers must be 16-byte aligned");
  PORL_CHECK_ARG(!(ema_beta > 0.f && hard_every > 0), "porl_adam_step_peer: choose Polyak or hard target update");
  if (pg->params[0] != nullptr) {           // two-shot: every rank updates one shard and stores it to all ranks
    ShardRank self;
    int rc = fill_shard_rank(ctx, self, p, pg, m, v, state_dev, target, n);
    if (rc != PORL_OK) return rc;
    const int blocks = sharded_blocks(ctx, n, pg->world);
    const double per = (double)n / pg->world;
    ProfScope prof(ctx, PORL_CAT_OPTIMIZER, per * (24.0 + 4.0 * pg->world + (target ? 4.0 : 0.0) + 4.0 * pg->world * (target ? 2.0 : 1.0)),
                   as_stream(stream));
    launch_pdl(ctx, adam_sharded_kernel, blocks, 256, 0, as_stream(stream), self, (const ShardRank*)nullptr, n, beta1, beta2,
               eps, ema_beta, (int)hard_every);
    PORL_LAUNCHED(ctx);
    return PORL_OK;
  }
  PeerArgs a{};
  a.world = pg->world;
  a.rank = pg->rank;
  for (int r = 0; r < pg->world; ++r) {
    PORL_CHECK_ARG(pg->bucket[r] && pg->pad[r] && aligned16(pg->bucket[r]), "porl_adam_step_peer: bucket/pad of rank %d", r);
    a.bucket[r] = pg->bucket[r];
    a.pad[r] = pg->pad[r];
  }
  a.epoch = pg->epoch;
  a.tail_out = pg->tail_out;
  a.tail_floats = pg->tail_floats;
  a.wait_ns = (unsigned long long)(pg->wait_ms > 0 ? pg->wait_ms : 20000) * 1000000ull;
  a.status = ctx->counters;
  int64_t blocks = ((n >> 2) + 255) / 256;
  const int64_t cap = (int64_t)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  ProfScope prof(ctx, PORL_CAT_OPTIMIZER, (double)n * ((target ? 32.0 : 24.0) + 4.0 * pg->world), as_stream(stream));
  launch_pdl(ctx, adam_kernel<true>, (int)blocks + 1 /* publisher block */, 256, 0, as_stream(stream), p, pg->bucket[pg->rank], m, v, n, state_dev, beta1,
                                                                beta2, eps, target, ema_beta, hard_every, a);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}

// ---- peer-mappable allocations (CUDA IPC) ----
extern "C" int porl_peer_alloc(porl_ctx_t* ctx, size_t bytes, void** out) {
  PORL_CHECK_ARG(ctx && out && bytes > 0, "porl_peer_alloc: null argument or zero size");
  PORL_CUDA(cudaSetDevice(ctx->device));
  void* ptr = nullptr;
  PORL_CUDA(cudaMalloc(&ptr, bytes));
  PORL_CUDA(cudaMemset(ptr, 0, bytes));
  PORL_CUDA(cudaDeviceSynchronize());
  *out = ptr;
  return PORL_OK;
}

extern "C" int porl_peer_free(porl_ctx_t* ctx, void* ptr) {
  PORL_CHECK_ARG(ctx, "porl_peer_free: null context");
  if (ptr) PORL_CUDA(cudaFree(ptr));
  return PORL_OK;
}

extern "C" int porl_peer_export(porl_ctx_t* ctx, void* ptr, uint8_t handle[64]) {
  PORL_CHECK_ARG(ctx && ptr && handle, "porl_peer_export: null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  cudaIpcMemHandle_t h;
  PORL_CUDA(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle, &h, 64);
  return PORL_OK;
}

extern "C" int porl_peer_open(porl_ctx_t* ctx, const uint8_t handle[64], void** out) {
  PORL_CHECK_ARG(ctx && handle && out, "porl_peer_open: null argument");
  PORL_CUDA(cudaSetDevice(ctx->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void* ptr = nullptr;
  const cudaError_t err = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
  if (err == cudaErrorPeerAccessUnsupported || err == cudaErrorInvalidDevice) {
    cudaGetLastError();
    set_error("porl_peer_open: the peer's device is not reachable from device %d (%s)", ctx->device, cudaGetErrorString(err));
    return PORL_ERR_UNSUPPORTED;
  }
  PORL_CUDA(err);
  *out = ptr;
  return PORL_OK;
}

extern "C" int porl_peer_close(porl_ctx_t* ctx, void* ptr) {
  PORL_CHECK_ARG(ctx, "porl_peer_close: null context");
  if (ptr) PORL_CUDA(cudaIpcCloseMemHandle(ptr));
  return PORL_OK;
}
