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

// learning rate after the scheduler step that follows update number e: torch CosineAnnealingLR.get_lr (recursive
// form, eta_min = 0), or the constant rate
__device__ __forceinline__ double adam_next_lr(long long e, double lr, double base_lr, long long t_max) {
  if (t_max <= 0) return lr;
  const double T = (double)t_max;
  if ((e - 1 - t_max) % (2 * t_max) == 0) return lr + base_lr * (1.0 - cos(M_PI / T)) / 2.0;
  return (1.0 + cos(M_PI * (double)e / T)) / (1.0 + cos(M_PI * (double)(e - 1) / T)) * lr;
}

// advance the optimizer clock by one step (scheduler.last_epoch after its step()).  One thread of the launch calls this
// after every block has read the clock.
__device__ __forceinline__ void adam_advance_clock(porl_opt_state_t* st) {
  const long long e = st->step + 1;
  if (st->t_max > 0) {
    st->lr = adam_next_lr(e, st->lr, st->base_lr, st->t_max);
    __threadfence();      // a reader that sees the new step also sees the new learning rate
  }
  st->step = e;
}

// ---- published scalars ---------------------------------------------------------------------------------------------
// adam_scalars() is a chain of dependent work on one thread (read the clock, two double-precision pow, a division, a
// square root: ~4 us) that every block of an optimizer launch has on its critical path before it may touch an element
// (measured: the 41-block Adam launch of the CQL net took 8 us, its 0.04 M parameters need < 1).  The scalars of update
// t + 1 depend only on the clock after update t, which is a pure function of the clock before it -- so block 0 of the
// launch of update t computes them on the side and publishes them in the optimizer state as ONE 16-byte record
// (slot = parity of t + 1; the blocks of this launch read the other slot).  A launch then needs one round trip: both
// records, step count and learning rate are loaded side by side and the record is verified against a check word that
// mixes every input the scalars depend on; a record that is missing, stale (the host rewrote the clock), foreign or
// torn fails the check and the block falls back to adam_scalars().  Both paths evaluate the same expressions: bitwise
// the same scalars.
struct AdamRecord {
  float step_size, bc2_sqrt;
  uint32_t t_lo, check;
};
static_assert(sizeof(AdamRecord) == 16, "one 16-byte access");

__device__ __forceinline__ uint32_t adam_record_check(float step_size, float bc2_sqrt, long long t, double lr, float beta1,
                                                      float beta2) {
  const unsigned long long lrb = (unsigned long long)__double_as_longlong(lr);
  uint32_t h = __float_as_uint(step_size) * 0x9E3779B1u;
  h = (h ^ __float_as_uint(bc2_sqrt)) * 0x85EBCA77u;
  h = (h ^ (uint32_t)t ^ (uint32_t)((unsigned long long)t >> 32)) * 0xC2B2AE3Du;
  h = (h ^ (uint32_t)lrb ^ (uint32_t)(lrb >> 32)) * 0x27D4EB2Fu;
  h = (h ^ __float_as_uint(beta1)) * 0x165667B1u;
  h = (h ^ __float_as_uint(beta2)) * 0x9E3779B1u;
  return h ^ (h >> 15) ^ 0xA5A5A5A5u;
}

// scalars of the update this launch applies (t = step + 1): from the published record when it verifies.  Two phases so
// that a caller can put other loads between issue and use: adam_clock_load() issues four independent loads,
// adam_scalars_from() consumes them.
struct AdamClock {
  uint4 r0, r1;
  long long t;        // update number this launch applies (step + 1)
  double lr;
};
__device__ __forceinline__ AdamClock adam_clock_load(porl_opt_state_t* st) {
  AdamClock c;
  c.r0 = __ldcv(reinterpret_cast<const uint4*>(&st->rec_[0][0]));
  c.r1 = __ldcv(reinterpret_cast<const uint4*>(&st->rec_[1][0]));
  c.t = *reinterpret_cast<const volatile long long*>(&st->step) + 1;
  c.lr = *reinterpret_cast<const volatile double*>(&st->lr);
  return c;
}
__device__ __forceinline__ AdamScalars adam_scalars_from(const AdamClock& c, porl_opt_state_t* st, float beta1, float beta2,
                                                         int hard_every) {
  const uint4 rr = (c.t & 1) ? c.r1 : c.r0;
  AdamRecord r;
  *reinterpret_cast<uint4*>(&r) = rr;
  if (r.t_lo == (uint32_t)c.t && r.check == adam_record_check(r.step_size, r.bc2_sqrt, c.t, c.lr, beta1, beta2)) {
    AdamScalars s;
    s.step_size = r.step_size;
    s.bc2_sqrt = r.bc2_sqrt;
    s.inv_bc2_sqrt = 0.f;
    s.do_hard = (hard_every > 0 && (c.t % hard_every) == 0) ? 1 : 0;
    return s;
  }
  return adam_scalars(st, beta1, beta2, hard_every);
}
__device__ __forceinline__ AdamScalars adam_scalars_fast(porl_opt_state_t* st, float beta1, float beta2, int hard_every,
                                                         long long* t_out = nullptr, double* lr_out = nullptr) {
  const AdamClock c = adam_clock_load(st);
  if (t_out) *t_out = c.t;
  if (lr_out) *lr_out = c.lr;
  return adam_scalars_from(c, st, beta1, beta2, hard_every);
}

// called by ONE thread of the launch that applies update t_cur (with the clock it read: t_cur = step + 1, lr_cur):
// publishes the scalars of update t_cur + 1, i.e. for the clock after this launch (adam_next_lr).  Reads nothing the
// clock advance writes, so it may run before or after it.
__device__ __forceinline__ void adam_publish_next(porl_opt_state_t* st, long long t_cur, double lr_cur, float beta1,
                                                  float beta2) {
  const double lr_next = adam_next_lr(t_cur, lr_cur, st->base_lr, st->t_max);
  const long long t = t_cur + 1;
  const double bc1 = 1.0 - pow((double)beta1, (double)t);       // the expressions of adam_scalars()
  const double bc2 = 1.0 - pow((double)beta2, (double)t);
  AdamRecord r;
  r.step_size = (float)(lr_next / bc1);
  r.bc2_sqrt = (float)sqrt(bc2);
  r.t_lo = (uint32_t)t;
  r.check = adam_record_check(r.step_size, r.bc2_sqrt, t, lr_next, beta1, beta2);
  __stwt(reinterpret_cast<uint4*>(&st->rec_[t & 1][0]), *reinterpret_cast<const uint4*>(&r));
}

}  // namespace porl
