"""Closed-form losses / gradients / optimizer maths used as kernel known-answer checks.

Test infrastructure.  Everything is dtype-generic torch on CPU (run it in float64 for the
yardstick).  tests/test_oracle_formulas.py checks each function against autograd of the
eager restatement in oracle.steps, which in turn is pinned to the reference by tests/golden.

Follows: agent/por.py:15-17,85-87,97-106; agent/sorl.py:102-109; agent/policy.py:18-23;
src/porl/train/cql_trainer.py:72-86,94-108; torch/optim/adam.py (single-tensor path);
torch/optim/lr_scheduler.py CosineAnnealingLR.get_lr (recursive form).
"""
from __future__ import annotations

import math

import torch

LOG_2PI = math.log(2.0 * math.pi)


def expectile_twin(v1, v2, next_v1, next_v2, r, done, gamma, tau):
    """TD target + twin expectile loss and d(loss)/dv_k  (por.py:85-87).

    loss = 1/2 * sum_k mean_i(|tau - 1[u_ik<0]| u_ik^2),  u_ik = target_i - v_ik
    dloss/dv_ik = -|tau - 1[u_ik<0]| u_ik / B
    """
    target = r + (1.0 - done) * gamma * torch.minimum(next_v1, next_v2)
    out, loss = [], 0.0
    for v in (v1, v2):
        u = target - v
        w = torch.abs(tau - (u < 0).to(u.dtype))
        loss = loss + (w * u * u).mean() / 2
        out.append(-(w * u) / u.numel())
    return target, loss, out[0], out[1]


def awr_gauss_nll(mean, log_std, x, target_v, v_new, alpha, mode):
    """Advantage-weighted diagonal-Gaussian NLL and its gradients  (por.py:97-106, sorl.py:102-109).

    mode "por": w = min(exp((target_v - v_new)/alpha), 100);  "sorl": w = min(exp(alpha*(target_v - v_new)), 100)
    nll_i = 1/2 sum_j z_ij^2 + sum_j ls_j + D/2 log 2pi, z = (x - mean)/exp(ls), ls = clamp(log_std, -5, 2)
    Returns loss, dmean (B,D), dlog_std (D,), weight, nll.
    """
    adv = target_v - v_new
    w = torch.exp(adv / alpha if mode == "por" else adv * alpha).clamp_max(100.0)
    ls = log_std.clamp(-5.0, 2.0)
    z = (x - mean) * torch.exp(-ls)
    nll = 0.5 * (z * z).sum(1) + ls.sum() + 0.5 * x.shape[1] * LOG_2PI
    n = x.shape[0]
    loss = (w * nll).mean()
    dmean = -(w[:, None] / n) * z * torch.exp(-ls)
    inside = ((log_std >= -5.0) & (log_std <= 2.0)).to(mean.dtype)
    dlog_std = ((w[:, None] / n) * (1.0 - z * z)).sum(0) * inside
    return loss, dmean, dlog_std, w, nll


def cql_head(q, q_next_tgt, actions, r, done, gamma, alpha):
    """CQL(H) discrete head: max-target TD MSE + alpha*(logsumexp - log A - Q(s,a))  (cql_trainer.py:72-108).

    dL/dQ_ia = (2/B)(Q_i,a_i - t_i) 1[a=a_i] + (alpha/B)(softmax(Q_i)_a - 1[a=a_i])
    Returns loss, td, penalty, dq, targets.
    """
    n, a = q.shape
    t = r + gamma * q_next_tgt.max(1).values * (1.0 - done)
    q_a = q.gather(1, actions[:, None]).squeeze(1)
    td = ((q_a - t) ** 2).mean()
    pen = (torch.logsumexp(q, 1) - math.log(a) - q_a).mean()
    onehot = torch.zeros_like(q).scatter_(1, actions[:, None], 1.0)
    dq = (2.0 / n) * (q_a - t)[:, None] * onehot + (alpha / n) * (torch.softmax(q, 1) - onehot)
    return td + alpha * pen, td, pen, dq, t


def adam_step(p, g, m, v, t, lr, beta1=0.9, beta2=0.999, eps=1e-8):
    """One torch.optim.Adam step (t = 1-based step count after increment); returns new (p, m, v)."""
    m = m + (g - m) * (1.0 - beta1)                       # exp_avg.lerp_(grad, 1-beta1)
    v = v * beta2 + (1.0 - beta2) * g * g
    step_size = lr / (1.0 - beta1 ** t)
    denom = v.sqrt() / math.sqrt(1.0 - beta2 ** t) + eps
    return p - step_size * (m / denom), m, v


def cosine_lr_sequence(base_lr, t_max, n, eta_min=0.0):
    """LR used at update 1..n by Adam under CosineAnnealingLR stepped once per update
    (torch's recursive get_lr incl. the restart branch); python-float arithmetic."""
    lrs, lr = [], base_lr
    for epoch in range(n):
        lrs.append(lr)
        e = epoch + 1  # last_epoch after the scheduler.step() that follows update `epoch`
        if (e - 1 - t_max) % (2 * t_max) == 0:
            lr = lr + (base_lr - eta_min) * (1 - math.cos(math.pi / t_max)) / 2
        else:
            lr = (1 + math.cos(math.pi * e / t_max)) / (1 + math.cos(math.pi * (e - 1) / t_max)) * (lr - eta_min) + eta_min
    return lrs
