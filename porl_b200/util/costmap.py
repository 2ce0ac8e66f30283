"""state2costmap, drop-in for util/costmap.py:7-64: (B, 362) scan + goal state -> (B, 3, 360, 256) raster
(channel 0: one-hot range bin per beam, rolled by 180 rows, bin 0 cleared; all channels: 5-pixel cross at the
goal's (bearing, distance) cell).  One kernel (csrc/encoder.cu, porl_costmap) writes the NHWC buffer whose
permuted view the reference returns (strides (276480, 1, 768, 3)).  Like the reference, the input is modified
in place: values > 8 become 0 (costmap.py:17).

The training path does not call this: the encoder evaluates the patch embedding sparsely from the scan and
never builds the 1.1 MB/sample raster (agent/fasternet.py of this package)."""
from __future__ import annotations

import torch

from .. import _C
from .._C import check, lib


def state2costmap(state: torch.Tensor) -> torch.Tensor:
    if state.dim() != 2 or state.shape[1] != 362:
        raise ValueError(f"expected (B, 362) states, got {tuple(state.shape)}")
    B = int(state.shape[0])
    ctx = _C.context(state.device)
    direct = state.dtype == torch.float32 and state.stride(1) == 1 and state.stride(0) >= 362
    buf = state if direct else state.to(torch.float32).contiguous()
    out = torch.empty(B, 360, 256, 3, dtype=torch.float32, device=state.device)
    if B:
        check(lib.porl_costmap(ctx.handle, buf.data_ptr(), buf.stride(0), B, out.data_ptr(), _C.stream_ptr()))
        if not direct:
            state.copy_(buf)
        flags = int(ctx.status_tensor().item())
        if flags & _C.FLAG_SCAN_RANGE:
            ctx.status_tensor().zero_()
            raise RuntimeError("state2costmap: scan value outside [0, 4.0001) u (8, inf) "
                               "(the reference's scatter_ raises here, util/costmap.py:29)")
    return out.permute(0, 3, 1, 2)
