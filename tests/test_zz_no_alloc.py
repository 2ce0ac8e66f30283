"""porl_expectile_twin_loss_noalloc (include/porl_b200.h): the no-allocation variant of K4 whose reduction scratch is a
file-scope static __device__ array.  No counterpart in the reference.
Called through the C ABI and compared with porl_expectile_twin_loss on the same inputs: same kernel, same order-fixed
reduction, so the outputs must be BITWISE equal (the tolerance of the K4 tests, 1e-6 relative, is asserted as well),
on two calls in a row (the second call sees the first call's scratch contents).  Collected last on
purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
LD = 4


def _inputs(B, seed):
    g = torch.Generator().manual_seed(seed)
    cols = lambda: torch.randn(B, LD, generator=g)
    v, nv = torch.stack([cols(), cols()]), torch.stack([cols(), cols()])
    r = torch.randn(B, generator=g)
    done = (torch.rand(B, generator=g) < 0.1).float()
    return [t.to(DEV).contiguous() for t in (v, nv, r, done)]


def _outputs(B):
    return (torch.full((B,), -7.0, device=DEV), torch.full((2, B, LD), -7.0, device=DEV), torch.full((8,), -7.0, device=DEV))


def _call_existing(ctx, _C, v, nv, r, done, B, tau, outs):
    tv, dv, loss = outs
    _C.check(_C.lib.porl_expectile_twin_loss(ctx.handle, v[0].data_ptr(), v[1].data_ptr(), nv[0].data_ptr(),
                                             nv[1].data_ptr(), LD, r.data_ptr(), done.data_ptr(), 0.99, tau, B, B,
                                             tv.data_ptr(), dv[0].data_ptr(), dv[1].data_ptr(), loss.data_ptr(),
                                             _C.stream_ptr()))


def _call_noalloc(_C, v, nv, r, done, B, tau, outs):
    tv, dv, loss = outs
    return _C.lib.porl_expectile_twin_loss_noalloc(v[0].data_ptr(), v[1].data_ptr(), nv[0].data_ptr(), nv[1].data_ptr(),
                                                   LD, r.data_ptr(), done.data_ptr(), 0.99, tau, B, B, tv.data_ptr(),
                                                   dv[0].data_ptr(), dv[1].data_ptr(), loss.data_ptr(), _C.stream_ptr())


def test_noalloc_expectile_matches_existing_entry_point_twice():
    from porl_b200 import _C
    torch.cuda.set_device(0)
    ctx = _C.context(torch.device(DEV))
    # two calls in a row with DIFFERENT sizes and data: the second runs over the first call's leftover partials
    for call, (B, tau, seed) in enumerate([(70_000, 0.9, 5), (1000, 0.7, 6)]):
        v, nv, r, done = _inputs(B, seed)
        want, got = _outputs(B), _outputs(B)
        _call_existing(ctx, _C, v, nv, r, done, B, tau, want)
        torch.cuda.synchronize()
        reserved, free = torch.cuda.memory_reserved(), torch.cuda.mem_get_info()[0]
        _C.check(_call_noalloc(_C, v, nv, r, done, B, tau, got))
        torch.cuda.synchronize()
        if call > 0:    # the first call per device resolves the symbol (may load the module); later calls hold steady
            assert torch.cuda.memory_reserved() == reserved and torch.cuda.mem_get_info()[0] == free
        for name, g_, w_ in zip(("target_v", "dv", "loss"), got, want):
            g_, w_ = (g_[:1], w_[:1]) if name == "loss" else (g_[..., 0], w_[..., 0])
            assert torch.isfinite(g_).all(), name
            torch.testing.assert_close(g_, w_, rtol=1e-6, atol=1e-6, msg=lambda m: f"call {call} {name}: {m}")
            assert torch.equal(g_, w_), f"call {call}: {name} is not bitwise equal to porl_expectile_twin_loss"
        assert torch.equal(got[1][..., 1:], torch.full_like(got[1][..., 1:], -7.0))      # padding columns untouched


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    from porl_b200 import _C
    torch.cuda.set_device(0)
    limit = 1 << 20                                     # PORL_NOALLOC_MAX_ROWS
    B = 1024
    v, nv, r, done = _inputs(B, 1)
    outs = _outputs(B)
    rc = _call_noalloc(_C, v, nv, r, done, limit + 1, 0.9, outs)       # rejected before any launch: nothing is read
    assert rc != 0
    msg = _C.lib.porl_last_error().decode()
    assert "PORL_NOALLOC_MAX_ROWS" in msg and str(limit) in msg, msg
    with pytest.raises(_C.PorlError):
        _C.check(rc)
    torch.cuda.synchronize()
    assert float(outs[2][0]) == -7.0                    # untouched
