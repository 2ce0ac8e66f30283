// Whole-update entry points of the C ABI (porl_sorl_step / porl_por_step): the launch sequence of one
// POR / SORL gradient update -- twin expectile value step, Polyak target update, advantage-weighted Gaussian
// log-likelihood policy step (agent/por.py:73-112, agent/sorl.py:78-128) -- issued from C on the caller's stream, so a
// host in any language runs an update with one call (and captures it into one CUDA graph).  No kernel lives here: the
// functions below only sequence porl_mlp_forward / porl_expectile_twin_loss / porl_mlp_backward / porl_adam_step[_peer]
// / porl_awr_gauss_nll over one caller-owned workspace.
#include <stdlib.h>

#include "common.cuh"

namespace {

struct StepWs {
  float *saved_t, *saved_v, *saved_v2, *saved_p, *out_t, *out_v, *out_v2, *out_p, *dv, *dz_p, *target_v;
  int64_t total;
};

int64_t up4(int64_t x) { return (x + 3) / 4 * 4; }

// carve the workspace (all blocks 16-byte aligned); base == nullptr only measures
StepWs carve(const porl_agent_step_t* a, int64_t B, float* base) {
  StepWs w;
  int64_t off = 0;
  auto take = [&](int64_t n) {
    float* p = base ? base + off : nullptr;
    off += up4(n);
    return p;
  };
  const int64_t sv = porl_mlp_saved_floats(a->value_mlp, B), sp = porl_mlp_saved_floats(a->policy_mlp, B);
  const int64_t ldv = porl_mlp_out_ld(a->value_mlp), ldp = porl_mlp_out_ld(a->policy_mlp);
  w.saved_t = take(sv);
  w.saved_v = take(sv);
  w.saved_v2 = take(sv);
  w.saved_p = take(sp);
  w.out_t = take(2 * B * ldv);
  w.out_v = take(2 * B * ldv);
  w.out_v2 = take(2 * B * ldv);
  w.dv = take(2 * B * ldv);
  w.out_p = take(B * ldp);
  w.dz_p = take(B * ldp);
  w.target_v = take(B);
  w.total = off;
  return w;
}

int group_step(porl_ctx_t* ctx, const porl_opt_group_t& g, float* target, float ema_beta, porl_stream_t st) {
  if (g.peers)
    return porl_adam_step_peer(ctx, g.params, g.peers, g.m, g.v, g.n, g.state, g.beta1, g.beta2, g.eps, target, ema_beta,
                               0, st);
  return porl_adam_step(ctx, g.params, g.grads, g.m, g.v, g.n, g.state, g.beta1, g.beta2, g.eps, target, ema_beta, 0, st);
}

int expectile_step(porl_ctx_t* ctx, const porl_agent_step_t* a, int64_t B, const float* x_obs, const float* x_next,
                   int64_t ld_x, const float* rewards, const float* dones, const float* reg_target, int64_t ld_target,
                   float* workspace, int64_t workspace_floats, int value_only, porl_stream_t st) {
  PORL_CHECK_ARG(ctx && a && a->value_mlp && a->policy_mlp && x_obs && x_next && rewards && dones && workspace,
                 "porl step: null argument");
  PORL_CHECK_ARG(B > 0 && a->world >= 1, "porl step: B=%lld world=%d", (long long)B, a->world);
  PORL_CHECK_ARG(a->value.params && a->value.grads && a->value.m && a->value.v && a->value.state && a->value.loss &&
                     a->target_params,
                 "porl step: value group incomplete");
  PORL_CHECK_ARG(value_only || (a->policy.params && a->policy.grads && a->policy.m && a->policy.v && a->policy.state &&
                                a->policy.loss && reg_target),
                 "porl step: policy group incomplete");
  StepWs w = carve(a, B, workspace);
  PORL_CHECK_ARG(workspace_floats >= w.total, "porl step: workspace holds %lld floats, %lld needed (porl_step_workspace_floats)",
                 (long long)workspace_floats, (long long)w.total);
  PORL_CHECK_ARG(aligned16(workspace), "porl step: workspace must be 16-byte aligned");
  const int64_t ldv = porl_mlp_out_ld(a->value_mlp), ldp = porl_mlp_out_ld(a->policy_mlp);
  const int64_t net = B * ldv;
  const int64_t b_norm = B * a->world;
  int rc;
  // 1. target TwinV(next) + online TwinV(obs)   (por.py:80-84, sorl.py:85-89)
  porl_mlp_call_t calls[2];
  const bool fused_head = porl_value_head_step_supported(a->value_mlp) != 0;
  calls[0] = porl_mlp_call_t{a->value_mlp, a->target_params, x_next, ld_x, w.saved_t, w.out_t, 1, fused_head ? PORL_MLP_SKIP_HEAD : 0};
  calls[1] = porl_mlp_call_t{a->value_mlp, a->value.params, x_obs, ld_x, w.saved_v, w.out_v, 0, fused_head ? PORL_MLP_SKIP_HEAD : 0};
  if ((rc = porl_mlp_forward(ctx, 2, calls, B, st)) != PORL_OK) return rc;
  if (fused_head) {
    // 2+3a. heads, TD target, twin expectile loss and the heads' backward in one kernel   (por.py:85-87)
    if ((rc = porl_value_head_step(ctx, &calls[0], &calls[1], B, rewards, dones, a->discount, a->tau, b_norm, w.target_v,
                                   a->value.grads, a->value.loss, st)) != PORL_OK)
      return rc;
    calls[1].pad_ = PORL_MLP_HEAD_DONE;
    if ((rc = porl_mlp_backward(ctx, &calls[1], B, nullptr, a->value.grads, st)) != PORL_OK) return rc;
  } else {
    // 2. TD target, twin expectile loss, dL/dv   (por.py:85-87)
    if ((rc = porl_expectile_twin_loss(ctx, w.out_v, w.out_v + net, w.out_t, w.out_t + net, ldv, rewards, dones,
                                       a->discount, a->tau, B, b_norm, w.target_v, w.dv, w.dv + net, a->value.loss,
                                       st)) != PORL_OK)
      return rc;
    // 3. backward of the online value nets into the group's bucket   (por.py:88-90)
    if ((rc = porl_mlp_backward(ctx, &calls[1], B, w.dv, a->value.grads, st)) != PORL_OK) return rc;
  }
  // 4. Adam + Polyak   (por.py:90-93)
  if ((rc = group_step(ctx, a->value, a->target_params, a->ema_beta, st)) != PORL_OK) return rc;
  if (value_only) return PORL_OK;
  // 5. online TwinV(obs) with the NEW weights + policy(obs)   (por.py:97-98,102)
  // PORL_FUSED_POLICY_HEAD=0: head forward, loss and head backward as three launches (kept for comparison)
  const char* const fe = getenv("PORL_FUSED_POLICY_HEAD");
  const bool fused_policy = porl_policy_head_step_supported(a->value_mlp, a->policy_mlp) != 0 && !(fe && atoi(fe) == 0);
  calls[0] = porl_mlp_call_t{a->value_mlp, a->value.params, x_obs, ld_x, w.saved_v2, w.out_v2, 1, fused_policy ? PORL_MLP_SKIP_HEAD : 0};
  calls[1] = porl_mlp_call_t{a->policy_mlp, a->policy.params, x_obs, ld_x, w.saved_p, w.out_p, 0, fused_policy ? PORL_MLP_SKIP_HEAD : 0};
  if ((rc = porl_mlp_forward(ctx, 2, calls, B, st)) != PORL_OK) return rc;
  if (fused_policy) {
    // 6+7a. value heads, advantage weight, Gaussian NLL and the policy head's backward in one kernel   (sorl.py:96-111)
    if ((rc = porl_policy_head_step(ctx, &calls[0], &calls[1], B, a->policy.params + a->log_std_offset, reg_target, ld_target,
                                    w.target_v, a->alpha, a->awr_mode, b_norm, a->policy.grads,
                                    a->policy.grads + a->log_std_offset, a->policy.loss, st)) != PORL_OK)
      return rc;
    calls[1].pad_ = PORL_MLP_HEAD_DONE;
    if ((rc = porl_mlp_backward(ctx, &calls[1], B, nullptr, a->policy.grads, st)) != PORL_OK) return rc;
    return group_step(ctx, a->policy, nullptr, 0.f, st);
  }
  // 6. advantage weight, Gaussian NLL, dL/dmean (tanh' folded), dL/dlog_std   (por.py:99-106, sorl.py:103-109)
  if ((rc = porl_awr_gauss_nll(ctx, w.out_p, ldp, a->policy.params + a->log_std_offset, reg_target, ld_target, w.target_v,
                               w.out_v2, w.out_v2 + net, ldv, a->alpha, a->awr_mode, a->tanh_out, B, a->policy_out, b_norm,
                               w.dz_p, a->policy.grads + a->log_std_offset, a->policy.loss, st)) != PORL_OK)
    return rc;
  // 7. policy backward, 8. Adam with the device-side cosine LR recursion   (por.py:107-110)
  if ((rc = porl_mlp_backward(ctx, &calls[1], B, w.dz_p, a->policy.grads, st)) != PORL_OK) return rc;
  return group_step(ctx, a->policy, nullptr, 0.f, st);
}

}  // namespace

extern "C" int64_t porl_step_workspace_floats(const porl_agent_step_t* agent, int64_t B) {
  if (!agent || !agent->value_mlp || !agent->policy_mlp || B <= 0) return -1;
  return carve(agent, B, nullptr).total;
}

extern "C" int porl_sorl_step(porl_ctx_t* ctx, const porl_agent_step_t* agent, int64_t B, const float* x_obs,
                              const float* x_next, int64_t ld_x, const float* actions, int64_t ld_act, const float* rewards,
                              const float* dones, float* workspace, int64_t workspace_floats, int32_t value_only,
                              porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  return expectile_step(ctx, agent, B, x_obs, x_next, ld_x, rewards, dones, actions, ld_act, workspace, workspace_floats,
                        value_only, stream);
}

extern "C" int porl_por_step(porl_ctx_t* ctx, const porl_agent_step_t* agent, int64_t B, const float* x_obs,
                             const float* x_next, int64_t ld_x, const float* next_obs, int64_t ld_next, const float* rewards,
                             const float* dones, float* workspace, int64_t workspace_floats, porl_stream_t stream) {
  PORL_ON_DEVICE(ctx);
  return expectile_step(ctx, agent, B, x_obs, x_next, ld_x, rewards, dones, next_obs, ld_next, workspace, workspace_floats,
                        0, stream);
}
