import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    def load(name):
        return dict(np.load(os.path.join(GOLDEN, name + ".npz")))

    return load


class _Recorder(list):
    pass


@pytest.fixture
def dry_porl(monkeypatch):
    """porl_b200 with every kernel-launching entry point replaced by a recording no-op and the device
    context faked on the CPU: exercises the HOST plumbing (parameter views, workspace sizing, call
    sequence, pointer arithmetic) without a GPU.  No arithmetic happens -- results are not checked here;
    parity lives in the -m gpu tests."""
    import ctypes
    import torch
    from porl_b200 import _C

    calls = _Recorder()
    launching = ["porl_gather_rows_f32", "porl_gather_transitions", "porl_gather_soa", "porl_pack_rows",
                 "porl_mlp_forward", "porl_mlp_backward", "porl_expectile_twin_loss", "porl_awr_gauss_nll",
                 "porl_cql_head", "porl_cql_fused_grad", "porl_cql_fused_step", "porl_adam_step", "porl_gemm", "porl_sorl_step", "porl_por_step"]
    for name in launching:
        monkeypatch.setattr(_C.lib, name, (lambda n: (lambda *a: calls.append((n, a)) or 0))(name))

    class FakeCtx:
        handle = ctypes.c_void_p(1)
        device_index = 0
        status_ptr = 0
        launches = 0
        gemm_mode = 0

        def status_tensor(self):
            return torch.zeros(1, dtype=torch.int32)

        def set_gemm_mode(self, m):
            self.gemm_mode = m

    ctx = FakeCtx()
    monkeypatch.setattr(_C, "context", lambda device: ctx)
    monkeypatch.setattr(_C, "stream_ptr", lambda: 0)
    monkeypatch.setattr(_C, "sync_stream", lambda: None)
    monkeypatch.setattr(_C, "pinned", lambda n, dtype: torch.zeros(n, dtype=dtype))
    return calls
