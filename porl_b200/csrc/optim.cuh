// Device-side pieces of the Adam step shared by the optimizer kernels (optim.cu) and by kernels that fold the optimizer
// into their own tail (cql_fused.cu): scalar set-up, the per-element update (torch/optim/adam.py, _single_tensor_adam)
// and the optimizer clock (step count + torch's recursive CosineAnnealingLR).
#pragma once
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

// advance the optimizer clock by one step: scheduler.last_epoch after its step(), torch CosineAnnealingLR.get_lr
// (recursive form, eta_min = 0).  One thread of the launch calls this after every element has been updated.
__device__ __forceinline__ void adam_advance_clock(porl_opt_state_t* st) {
  const long long e = st->step + 1;
  if (st->t_max > 0) {
    const double T = (double)st->t_max;
    double lr = st->lr;
    if ((e - 1 - st->t_max) % (2 * st->t_max) == 0) lr = lr + st->base_lr * (1.0 - cos(M_PI / T)) / 2.0;
    else lr = (1.0 + cos(M_PI * (double)e / T)) / (1.0 + cos(M_PI * (double)(e - 1) / T)) * lr;
    st->lr = lr;
    __threadfence();      // a reader that sees the new step also sees the new learning rate
  }
  st->step = e;
}

}  // namespace porl
