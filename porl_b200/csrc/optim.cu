// K7: fused Adam step over a flat parameter buffer, with the Polyak / hard target update folded in
// and the optimizer clock (step count, cosine-annealed LR) kept on the device.
//
// HBM-bound: 28 B/param (read p,g,m,v; write p,m,v) + 8 B/param when a target net rides along.
// float4 accesses, grid-stride, one pass.  The scalar maths follows torch/optim/adam.py
// (_single_tensor_adam): bias corrections and step size in double, per-element work in fp32.
#include <math.h>

#include "common.cuh"

namespace porl {

struct AdamScalars {
  float step_size;    // lr / (1 - beta1^t)
  float inv_bc2_sqrt; // 1 / sqrt(1 - beta2^t)   (torch divides by bias_correction2_sqrt)
  float bc2_sqrt;
  int do_hard;        // copy p -> target this step
};

__device__ __forceinline__ AdamScalars adam_scalars(const porl_opt_state_t* st, float beta1, float beta2,
                                                    int hard_every) {
  const long long t = st->step + 1;
  const double bc1 = 1.0 - pow((double)beta1, (double)t);
  const double bc2 = 1.0 - pow((double)beta2, (double)t);
  AdamScalars s;
  s.step_size = (float)(st->lr / bc1);
  s.bc2_sqrt = (float)sqrt(bc2);
  s.inv_bc2_sqrt = 0.f;
  s.do_hard = (hard_every > 0 && (t % hard_every) == 0) ? 1 : 0;
  return s;
}

__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, const AdamScalars& s,
                                          float beta1, float beta2, float eps) {
  m = m + (g - m) * (1.f - beta1);                    // exp_avg.lerp_(grad, 1 - beta1)
  v = v * beta2 + (1.f - beta2) * g * g;              // mul_(beta2).addcmul_(g, g, 1 - beta2)
  const float denom = sqrtf(v) / s.bc2_sqrt + eps;    // (exp_avg_sq.sqrt() / bc2_sqrt).add_(eps)
  p = p - s.step_size * (m / denom);                  // addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
            float* __restrict__ v, int64_t n, porl_opt_state_t* st, float beta1, float beta2, float eps,
            float* __restrict__ target, float ema_beta, int hard_every) {
  __shared__ AdamScalars sh;
  if (threadIdx.x == 0) sh = adam_scalars(st, beta1, beta2, hard_every);
  __syncthreads();
  const AdamScalars s = sh;
  const bool ema = (target != nullptr) && ema_beta > 0.f;
  const bool hard = (target != nullptr) && s.do_hard;
  const float keep = 1.f - ema_beta;

  const int64_t n4 = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float4* p4 = reinterpret_cast<float4*>(p);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  float4* t4 = reinterpret_cast<float4*>(target);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pp = p4[i], mm = m4[i], vv = v4[i];
    const float4 gg = g4[i];
    adam_elem(pp.x, gg.x, mm.x, vv.x, s, beta1, beta2, eps);
    adam_elem(pp.y, gg.y, mm.y, vv.y, s, beta1, beta2, eps);
    adam_elem(pp.z, gg.z, mm.z, vv.z, s, beta1, beta2, eps);
    adam_elem(pp.w, gg.w, mm.w, vv.w, s, beta1, beta2, eps);
    p4[i] = pp; m4[i] = mm; v4[i] = vv;
    if (ema) {
      float4 tt = t4[i];
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
  for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float pp = p[i], mm = m[i], vv = v[i];
    adam_elem(pp, g[i], mm, vv, s, beta1, beta2, eps);
    p[i] = pp; m[i] = mm; v[i] = vv;
    if (ema) target[i] = fmaf(ema_beta, pp, target[i] * keep);
    else if (hard) target[i] = pp;
  }

  // the last block to finish advances the optimizer clock
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t ticket = atomicAdd(&st->blocks_done, 1u);
    if (ticket == gridDim.x - 1) {
      st->blocks_done = 0;
      const long long e = st->step + 1;   // scheduler.last_epoch after its step()
      if (st->t_max > 0) {
        // torch CosineAnnealingLR.get_lr (recursive form, eta_min = 0)
        const double T = (double)st->t_max;
        double lr = st->lr;
        if ((e - 1 - st->t_max) % (2 * st->t_max) == 0) {
          lr = lr + st->base_lr * (1.0 - cos(M_PI / T)) / 2.0;
        } else {
          lr = (1.0 + cos(M_PI * (double)e / T)) / (1.0 + cos(M_PI * (double)(e - 1) / T)) * lr;
        }
        st->lr = lr;
      }
      st->step = e;
      __threadfence();
    }
  }
}

}  // namespace porl

using namespace porl;

extern "C" int porl_adam_step(porl_ctx_t* ctx, float* p, const float* g, float* m, float* v, int64_t n,
                              porl_opt_state_t* state_dev, float beta1, float beta2, float eps,
                              float* target, float ema_beta, int32_t hard_every, porl_stream_t stream) {
  PORL_CHECK_ARG(ctx && p && g && m && v && state_dev, "porl_adam_step: null argument");
  PORL_CHECK_ARG(n > 0, "porl_adam_step: n=%lld", (long long)n);
  PORL_CHECK_ARG(aligned16(p) && aligned16(g) && aligned16(m) && aligned16(v) && (!target || aligned16(target)),
                 "porl_adam_step: buffers must be 16-byte aligned");
  PORL_CHECK_ARG(!(ema_beta > 0.f && hard_every > 0), "porl_adam_step: choose Polyak or hard target update");
  int64_t blocks = ((n >> 2) + 255) / 256;
  const int64_t cap = (int64_t)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  ProfScope prof(ctx, PORL_CAT_OPTIMIZER, (double)n * (target ? 36.0 : 28.0), as_stream(stream));
  adam_kernel<<<(int)blocks, 256, 0, as_stream(stream)>>>(p, g, m, v, n, state_dev, beta1, beta2, eps,
                                                          target, ema_beta, hard_every);
  PORL_LAUNCHED(ctx);
  return PORL_OK;
}
