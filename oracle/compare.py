"""Parity comparators shared by tests/ and __graft_entry__.smoke().  Test infrastructure.

Parameter tolerance after n Adam steps (lr = learning rate), from the reference's own fp32-vs-fp64
yardstick: Adam's m/(sqrt(v)+eps) turns rounding noise on a near-zero gradient
into a step of up to +-lr (sign flip: 2*lr), so the elementwise bound is stated in units of lr,
while the bulk of each tensor must agree to fp32 rounding:
  * per-tensor relative L2 distance          <= 5e-5
  * elements outside (rtol 5e-5, atol 2e-7)  <= 0.01 % of the tensor (at least 2 elements allowed)
  * every element                            |got - ref| <= 2 * lr * n_steps
"""
from __future__ import annotations

import numpy as np
import torch


def _np(x):
    return x.detach().double().cpu().numpy() if torch.is_tensor(x) else np.asarray(x, dtype=np.float64)


def rel_l2(a, b) -> float:
    a, b = _np(a).ravel(), _np(b).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def assert_params_close(got, ref, *, lr: float, n_steps: int = 1, name: str = "") -> None:
    g, r = _np(got), _np(ref)
    assert g.shape == r.shape, (name, g.shape, r.shape)
    d = np.abs(g - r)
    assert float(d.max(initial=0.0)) <= 2.0 * lr * n_steps + 1e-12, (name, "max abs", float(d.max()))
    if np.linalg.norm(r) > 0:
        assert rel_l2(g, r) <= 5e-5, (name, "rel-L2", rel_l2(g, r))
    bad = int((d > 2e-7 + 5e-5 * np.abs(r)).sum())
    assert bad <= max(2, int(1e-4 * g.size)), (name, "outliers", bad, g.size)
