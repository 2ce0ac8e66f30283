"""ctypes binding of libporl_b200.so (the C ABI declared in include/porl_b200.h).

No CPU fallback: importing this module without the built library raises, and creating a
context without an sm_100 device raises PorlError.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libporl_b200.so")

PAD = 32
MAX_LAYERS = 6
MAX_CALLS = 4

FLAG_NONFINITE_LOSS = 1
FLAG_NONPOSITIVE_NLL = 2
FLAG_BAD_INDEX = 4
FLAG_BAD_ACTION = 8
FLAG_SCAN_RANGE = 16

AWR_POR, AWR_SORL = 0, 1
PARAM_WEIGHT, PARAM_BIAS, PARAM_LN_WEIGHT, PARAM_LN_BIAS = 0, 1, 2, 3


class PorlError(RuntimeError):
    pass


class MlpDesc(C.Structure):
    _fields_ = [("n_nets", C.c_int32), ("n_layers", C.c_int32), ("dims", C.c_int32 * (MAX_LAYERS + 1)),
                ("layer_norm", C.c_int32), ("out_act", C.c_int32)]


class MlpCall(C.Structure):
    _fields_ = [("mlp", C.c_void_p), ("params", C.c_void_p), ("x", C.c_void_p), ("ldx", C.c_int64),
                ("saved", C.c_void_p), ("out", C.c_void_p), ("inference", C.c_int32), ("pad_", C.c_int32)]


class OptState(C.Structure):
    _fields_ = [("step", C.c_int64), ("lr", C.c_double), ("base_lr", C.c_double), ("t_max", C.c_int64),
                ("blocks_done", C.c_uint32), ("pad_", C.c_uint32), ("pad2_", C.c_uint32 * 2),
                ("rec_", (C.c_uint32 * 4) * 2)]


MAX_PEERS = 8
PEER_PAD_WORDS = 2 * MAX_PEERS
FLAG_PEER_TIMEOUT = 32


class PeerGroup(C.Structure):                      # porl_peer_group_t
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("bucket", C.c_void_p * MAX_PEERS),
                ("pad", C.c_void_p * MAX_PEERS), ("epoch", C.c_void_p), ("tail_out", C.c_void_p),
                ("tail_floats", C.c_int32), ("wait_ms", C.c_int32),
                ("params", C.c_void_p * MAX_PEERS), ("target", C.c_void_p * MAX_PEERS)]   # two-shot mode when params[0]


class PeerRank(C.Structure):                       # porl_peer_rank_t (single-GPU emulation of several ranks, tests)
    _fields_ = [("p", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p), ("target", C.c_void_p), ("state", C.c_void_p),
                ("peers", C.POINTER(PeerGroup))]


class OptGroup(C.Structure):                       # porl_opt_group_t
    _fields_ = [("params", C.c_void_p), ("grads", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p), ("n", C.c_int64),
                ("state", C.c_void_p), ("loss", C.c_void_p), ("beta1", C.c_float), ("beta2", C.c_float),
                ("eps", C.c_float), ("pad_", C.c_int32), ("peers", C.POINTER(PeerGroup))]


class AgentStep(C.Structure):                      # porl_agent_step_t
    _fields_ = [("value_mlp", C.c_void_p), ("policy_mlp", C.c_void_p), ("value", OptGroup), ("policy", OptGroup),
                ("target_params", C.c_void_p), ("log_std_offset", C.c_int64), ("discount", C.c_float),
                ("tau", C.c_float), ("alpha", C.c_float), ("ema_beta", C.c_float), ("awr_mode", C.c_int32),
                ("tanh_out", C.c_int32), ("policy_out", C.c_int32), ("world", C.c_int32)]


CATEGORIES = ("contraction", "gather", "loss", "optimizer", "norm", "transpose", "head", "encoder", "batchnorm", "pconv")


class Profile(C.Structure):
    _fields_ = [("launches", C.c_int64 * len(CATEGORIES)), ("ms", C.c_double * len(CATEGORIES)), ("work", C.c_double * len(CATEGORIES))]


_p, _i32, _i64, _f = C.c_void_p, C.c_int32, C.c_int64, C.c_float


class BnParams(C.Structure):                       # porl_bn_t
    _fields_ = [("weight", _p), ("bias", _p), ("running_mean", _p), ("running_var", _p)]


class FnBlock(C.Structure):                        # porl_fn_block_t
    _fields_ = [("pconv_w", _p), ("w1", _p), ("bn", BnParams), ("w2", _p), ("keep", _p)]


class FasterNetParams(C.Structure):                # porl_fasternet_t
    _fields_ = [("pe_w", _p), ("pe_bn", BnParams), ("block", FnBlock * 3), ("merge_w", _p), ("merge_bn", BnParams),
                ("head_conv_w", _p), ("head_w", _p), ("head_b", _p), ("feature_dim", _i32), ("world", _i32),
                ("stats_allreduce", _p), ("stats_user", _p), ("stats_buf", _p)]


ALLREDUCE_F64_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p)   # porl_allreduce_f64_fn


PM_MAX_STAGES, PM_MAX_BLOCKS = 4, 4


class PmConv(C.Structure):                         # porl_pm_conv_t
    _fields_ = [("weight", C.c_void_p), ("bias", C.c_void_p), ("bn", BnParams)]


class PmRes(C.Structure):                          # porl_pm_res_t
    _fields_ = [("net1", PmConv), ("net2", PmConv), ("c_mid", C.c_int32), ("pad_", C.c_int32)]


class PmStage(C.Structure):                        # porl_pm_stage_t
    _fields_ = [("anchors", C.c_int32), ("k", C.c_int32), ("c_in", C.c_int32), ("c_out", C.c_int32),
                ("affine_alpha", C.c_void_p), ("affine_beta", C.c_void_p), ("transfer", PmConv),
                ("n_pre", C.c_int32), ("n_pos", C.c_int32), ("pre", PmRes * PM_MAX_BLOCKS), ("pos", PmRes * PM_MAX_BLOCKS)]


class PointMlpParams(C.Structure):                 # porl_pointmlp_t
    _fields_ = [("points", C.c_int32), ("embed_dim", C.c_int32), ("n_stages", C.c_int32), ("use_xyz", C.c_int32),
                ("normalize", C.c_int32), ("class_num", C.c_int32), ("pad0_", C.c_int32), ("pad1_", C.c_int32),
                ("embedding", PmConv), ("stage", PmStage * PM_MAX_STAGES),
                ("fc1_w", C.c_void_p), ("fc1_b", C.c_void_p), ("bn1", BnParams),
                ("fc2_w", C.c_void_p), ("fc2_b", C.c_void_p), ("bn2", BnParams),
                ("fc3_w", C.c_void_p), ("fc3_b", C.c_void_p), ("drop1", C.c_void_p), ("drop2", C.c_void_p)]


_SIGNATURES = {
    "porl_version": (C.c_int, []),
    "porl_last_error": (C.c_char_p, []),
    "porl_ctx_create": (C.c_int, [C.POINTER(_p), C.c_int]),
    "porl_ctx_destroy": (C.c_int, [_p]),
    "porl_launch_count": (_i64, [_p]),
    "porl_tc_launch_count": (_i64, [_p]),
    "porl_ctx_status_ptr": (C.c_int, [_p, C.POINTER(_p)]),
    "porl_ctx_set_gemm_mode": (C.c_int, [_p, C.c_int]),
    "porl_ctx_get_gemm_mode": (C.c_int, [_p]),
    "porl_profile_enable": (C.c_int, [_p, C.c_int]),
    "porl_profile_read": (C.c_int, [_p, C.POINTER(Profile)]),
    "porl_gather_rows_f32": (C.c_int, [_p, _p, _i64, _i64, _i64, _p, _i64, _p, _i64, _p]),
    "porl_gather_transitions": (C.c_int, [_p, _p, _i64, _i64, _i32, _i32, _p, _i64, _p, _p, _i64, _p, _p, _p, _i64, _p]),
    "porl_gather_soa": (C.c_int, [_p, _p, _p, _p, _p, _p, _i64, _i32, _i64, _p, _i64, _p, _p, _i64, _p, _p, _p, _p]),
    "porl_sample_permutation": (C.c_int, [_p, _i64, C.c_uint64, C.c_uint64, _i64, _i64, _p, _p]),
    "porl_pack_rows": (C.c_int, [_p, _p, _i64, _i64, _i32, _p, _i64, _p]),
    "porl_mlp_create": (C.c_int, [_p, C.POINTER(MlpDesc), C.POINTER(_p)]),
    "porl_mlp_destroy": (C.c_int, [_p]),
    "porl_mlp_param_floats": (_i64, [_p]),
    "porl_mlp_param_layout": (C.c_int, [_p, _i32, _i32, _i32, C.POINTER(_i64), C.POINTER(_i64)]),
    "porl_mlp_saved_floats": (_i64, [_p, _i64]),
    "porl_mlp_out_ld": (_i64, [_p]),
    "porl_mlp_forward": (C.c_int, [_p, _i32, C.POINTER(MlpCall), _i64, _p]),
    "porl_mlp_backward": (C.c_int, [_p, C.POINTER(MlpCall), _i64, _p, _p, _p]),
    "porl_value_head_step_supported": (C.c_int, [_p]),
    "porl_value_head_step": (C.c_int, [_p, C.POINTER(MlpCall), C.POINTER(MlpCall), _i64, _p, _p, _f, _f, _i64, _p, _p, _p, _p]),
    "porl_policy_head_step_supported": (C.c_int, [_p, _p]),
    "porl_policy_head_step": (C.c_int, [_p, C.POINTER(MlpCall), C.POINTER(MlpCall), _i64, _p, _p, _i64, _p, _f, _i32, _i64,
                                        _p, _p, _p, _p]),
    "porl_expectile_twin_loss": (C.c_int, [_p, _p, _p, _p, _p, _i64, _p, _p, _f, _f, _i64, _i64, _p, _p, _p, _p, _p]),
    "porl_expectile_twin_loss_noalloc": (C.c_int, [_p, _p, _p, _p, _i64, _p, _p, _f, _f, _i64, _i64, _p, _p, _p, _p, _p]),
    "porl_awr_gauss_nll": (C.c_int, [_p, _p, _i64, _p, _p, _i64, _p, _p, _p, _i64, _f, _i32, _i32, _i64, _i32, _i64,
                                     _p, _p, _p, _p]),
    "porl_cql_head": (C.c_int, [_p, _p, _p, _i64, _p, _p, _p, _f, _f, _i64, _i32, _i64, _p, _p, _p]),
    "porl_cql_fused_supported": (C.c_int, [_p]),
    "porl_cql_fused_grad": (C.c_int, [_p, _p, _p, _p, _p, _p, _i64, _p, _p, _p, _p, _i64, _f, _f, _i64, _i64, _p, _p, _p]),
    "porl_cql_fused_step": (C.c_int, [_p, _p, _p, _p, _p, _p, _i64, _p, _p, _p, _p, _i64, _f, _f, _i64, _p, _p, _p, _p, _p,
                                      _f, _f, _f, _i32, _p]),
    "porl_mlp_desc": (C.c_int, [_p, _p]),
    "porl_adam_step": (C.c_int, [_p, _p, _p, _p, _p, _i64, _p, _f, _f, _f, _p, _f, _i32, _p]),
    "porl_pointmlp_workspace_floats": (_i64, [C.POINTER(PointMlpParams), _i64]),
    "porl_pointmlp_forward": (C.c_int, [_p, C.POINTER(PointMlpParams), _p, _p, _i64, _p, _i64, _p, _i64, _p]),
    "porl_fps": (C.c_int, [_p, _p, _i64, _i32, _i32, _p, _p, _p]),
    "porl_knn": (C.c_int, [_p, _p, _p, _i64, _i32, _i32, _i32, _p, _p]),
    "porl_scan_to_points": (C.c_int, [_p, _p, _i64, _i64, _p, _p]),
    "porl_adam_step_peer": (C.c_int, [_p, _p, C.POINTER(PeerGroup), _p, _p, _i64, _p, _f, _f, _f, _p, _f, _i32, _p]),
    "porl_adam_step_peer_emulated": (C.c_int, [_p, _i32, C.POINTER(PeerRank), _i64, _f, _f, _f, _f, _i32, _p]),
    "porl_peer_alloc": (C.c_int, [_p, C.c_size_t, C.POINTER(C.c_void_p)]),
    "porl_peer_free": (C.c_int, [_p, _p]),
    "porl_peer_export": (C.c_int, [_p, _p, C.c_char_p]),
    "porl_peer_open": (C.c_int, [_p, C.c_char_p, C.POINTER(C.c_void_p)]),
    "porl_peer_close": (C.c_int, [_p, _p]),
    "porl_step_workspace_floats": (_i64, [C.POINTER(AgentStep), _i64]),
    "porl_sorl_step": (C.c_int, [_p, C.POINTER(AgentStep), _i64, _p, _p, _i64, _p, _i64, _p, _p, _p, _i64, _i32, _p]),
    "porl_por_step": (C.c_int, [_p, C.POINTER(AgentStep), _i64, _p, _p, _i64, _p, _i64, _p, _p, _p, _i64, _p]),
    "porl_gemm": (C.c_int, [_p, _i32, _i32, _i64, _i64, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _i32, _p, _i64,
                            _i64, _p, _p, _i64, _p]),
    "porl_transpose": (C.c_int, [_p, _p, _i64, _i64, _i64, _p, _i64, _p]),
    "porl_costmap": (C.c_int, [_p, _p, _i64, _i64, _p, _p]),
    "porl_fasternet_workspace_floats": (_i64, [_i64]),
    "porl_fasternet_forward": (C.c_int, [_p, C.POINTER(FasterNetParams), _p, _i64, _i64, _p, _i64, _p, _i64, _p]),
    "porl_ctx_set_tc_chunk": (C.c_int, [_p, C.c_int]),
    "porl_ctx_set_tc_terms": (C.c_int, [_p, C.c_int]),
    "porl_ctx_set_tc_phases": (C.c_int, [_p, C.c_int]),
    "porl_ctx_set_tc_debug": (C.c_int, [_p, _p]),
    "porl_ctx_set_tc_variant": (C.c_int, [_p, C.c_int]),
    "porl_ctx_set_pdl": (C.c_int, [_p, C.c_int]),
}

EXPORTS = tuple(_SIGNATURES)


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m porl_b200.build` "
            "(porl_b200 has no CPU or PyTorch fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype, fn.argtypes = res, args
    return lib


lib = _load()


def check(status: int) -> None:
    if status != 0:
        raise PorlError(f"libporl_b200 error {status}: {lib.porl_last_error().decode(errors='replace')}")


_CONTEXTS: dict[int, "Context"] = {}


class Context:
    """One per device; owns the library's internal workspaces."""

    def __init__(self, device_index: int):
        h = _p()
        check(lib.porl_ctx_create(C.byref(h), device_index))
        self.handle = h
        self.device_index = device_index
        sp = _p()
        check(lib.porl_ctx_status_ptr(self.handle, C.byref(sp)))
        self.status_ptr = sp.value
        self.epoch = 0            # bumped by every option setter: CUDA graphs captured under older options are stale
        self.profiling = False

    def status_tensor(self):
        """Zero-copy int32 torch view of the device status word (PORL_FLAG_* bits)."""
        import torch

        class _Word:
            __cuda_array_interface__ = {"shape": (1,), "typestr": "<i4", "data": (self.status_ptr, False),
                                        "version": 3}

        return torch.as_tensor(_Word(), device=torch.device("cuda", self.device_index))

    @property
    def launches(self) -> int:
        return int(lib.porl_launch_count(self.handle))

    def profile(self, enable: bool) -> None:
        check(lib.porl_profile_enable(self.handle, int(enable)))
        self.profiling, self.epoch = bool(enable), self.epoch + 1

    def profile_read(self) -> dict:
        """{category: {"launches", "ms", "work"}} of the launches recorded since profile(True)."""
        out = Profile()
        check(lib.porl_profile_read(self.handle, C.byref(out)))
        return {c: {"launches": int(out.launches[i]), "ms": float(out.ms[i]), "work": float(out.work[i])}
                for i, c in enumerate(CATEGORIES)}

    @property
    def tc_launches(self) -> int:
        return int(lib.porl_tc_launch_count(self.handle))

    def set_gemm_mode(self, mode: int) -> None:
        check(lib.porl_ctx_set_gemm_mode(self.handle, mode))
        self.epoch += 1

    def set_tc_chunk(self, k_blocks: int) -> None:
        check(lib.porl_ctx_set_tc_chunk(self.handle, k_blocks))
        self.epoch += 1

    def set_tc_phases(self, enable: bool) -> None:
        check(lib.porl_ctx_set_tc_phases(self.handle, int(bool(enable))))
        self.epoch += 1

    def set_tc_terms(self, terms: int) -> None:
        check(lib.porl_ctx_set_tc_terms(self.handle, terms))
        self.epoch += 1

    def set_tc_variant(self, a_in_tmem: int) -> None:
        check(lib.porl_ctx_set_tc_variant(self.handle, a_in_tmem))
        self.epoch += 1

    def set_pdl(self, enable: bool) -> None:
        check(lib.porl_ctx_set_pdl(self.handle, int(bool(enable))))
        self.epoch += 1

    def set_tc_debug(self, ptr) -> None:
        check(lib.porl_ctx_set_tc_debug(self.handle, ptr))
        self.epoch += 1

    @property
    def gemm_mode(self) -> int:
        return int(lib.porl_ctx_get_gemm_mode(self.handle))


def context(device) -> Context:
    """Context for a torch.device / index (created on first use)."""
    import torch

    dev = torch.device(device)
    if dev.type != "cuda":
        raise PorlError(f"porl_b200 runs on CUDA (sm_100a) devices only, got device '{dev}'; there is no CPU fallback")
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    if idx not in _CONTEXTS:
        _CONTEXTS[idx] = Context(idx)
    return _CONTEXTS[idx]


def stream_ptr() -> int:
    import torch

    return torch.cuda.current_stream().cuda_stream


def sync_stream() -> None:
    import torch

    torch.cuda.current_stream().synchronize()


def pinned(n: int, dtype):
    import torch

    return torch.zeros(n, dtype=dtype).pin_memory()
