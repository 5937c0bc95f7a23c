"""ctypes binding of libb200grpo.so (the C ABI declared in include/b200grpo.h).

The library is the product: there is no Python/torch fallback.  ``lib()`` raises if the
shared object has not been built (run ``python -c 'import __graft_entry__ as g; g.build()'``
or ``make -C pipelinerl_swe_b200/csrc``); every call raises ``B200GrpoError`` with the
library's own message when the C side reports an error (e.g. no sm_100 device).
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200GRPO_LIB_PATH") or os.path.join(_HERE, "libb200grpo.so")   # the override serves same-box A/B runs of two builds
M_COUNT = 40  # B200GRPO_M_COUNT

# metric slots (enum in include/b200grpo.h)
(M_LOSS, M_N_MASKED, M_N_SEQ, M_NUM_NANS, M_REWARD, M_ENTROPY, M_OLD_LP, M_NEW_LP, M_REF_LP, M_ADVANTAGE, M_KL, M_POLICY,
 M_SURR1, M_SURR2, M_RATIO, M_RATIO_SUM, M_RATIO_SQ_SUM, M_RATIO_REF_NEW, M_RATIO_REF_OLD, M_CLAMP_REF_NEW,
 M_CLAMP_NEW_OLD, M_TOKEN_WEIGHT, M_VALUE_MEAN, M_VALUE_LOSS, M_VALUE_MSE, M_MAX_REWARD, M_MIN_REWARD, M_MAX_ADV,
 M_MIN_ADV, M_MAX_KL, M_MIN_KL, M_MAX_TW, M_MIN_TW, M_MAX_VALUE, M_MIN_VALUE, M_FLAGS) = range(36)
F_NEW_LP_NONFINITE, F_REF_NEW_NONFINITE, F_KL_NONFINITE, F_LOSS_NONFINITE, F_GROUP_TOKENS_LE0 = 1, 2, 4, 8, 16
BWD_DLOGITS, BWD_DH, BWD_DW = 1, 2, 4                                      # B200GRPO_BWD_*
STAT_SUM, STAT_MIN, STAT_MAX = 0, 1, 2                                    # B200GRPO_STAT_*
STAT_KL_COEF, STAT_ENTROPY_COEF, STAT_INPUT_SIZE = -1, -2, -3
STATS_MAX = 64


class B200GrpoError(RuntimeError):
    pass


class LossCfg(C.Structure):
    """b200grpo_loss_cfg"""
    _fields_ = [
        ("policy_loss", C.c_int32), ("use_advantages", C.c_int32), ("relu_log_p_weights", C.c_int32),
        ("group_normalization", C.c_int32), ("overlong_filtering", C.c_int32), ("is_packed", C.c_int32),
        ("has_values", C.c_int32), ("batch_size", C.c_int32),
        ("epsilon", C.c_float), ("clamp_log_ratio_ref_new", C.c_float), ("kl_coef", C.c_float),
        ("entropy_coef", C.c_float), ("value_loss_coef", C.c_float), ("reserved", C.c_float),
    ]


class HostBatch(C.Structure):
    """b200grpo_host_batch"""
    _fields_ = [(n, C.c_void_p) for n in (
        "hidden_bf16", "input_ids", "labels", "position_ids", "rewards", "advantages", "ref_logprobs", "old_logprobs",
        "group_tokens", "num_labels", "overflow")]


class PackedOut(C.Structure):
    """b200grpo_packed_out"""
    _fields_ = [(n, C.c_void_p) for n in (
        "input_ids", "labels", "attention_mask", "position_ids", "rewards", "advantages", "ref_logprobs", "old_logprobs",
        "group_tokens", "num_labels", "overflow")]


_p, _i, _i64, _f, _sz, _d = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t, C.c_double

# name -> (restype, argtypes); mirrors include/b200grpo.h one to one
SIGNATURES = {
    "b200grpo_last_error": (C.c_char_p, []),
    "b200grpo_abi_version": (_i, []),
    "b200grpo_set_tunable": (_i, [C.c_char_p, _i]),
    "b200grpo_launch_count": (C.c_uint64, []),
    "b200grpo_lmhead_fwd_workspace": (_sz, [_i, _i, _i]),
    "b200grpo_lmhead_saved_bytes": (_sz, [_i, _i, _i]),
    "b200grpo_lmhead_fwd": (_i, [_p, _i64, _p, _i64, _p, _i, _i, _i, _i64, _f, _p, _p, _p, _sz, _p]),
    "b200grpo_lmhead_fwd_multicast": (_i, [_p, _i64, _p, _i64, _p, _i, _i, _i, _i64, _f, _p, _p, _p, _sz, _p]),
    "b200grpo_mc_allreduce_bf16": (_i, [_p, _i64, _i, _i, _i, _p]),
    "b200grpo_lmhead_finalize": (_i, [_p, _i, _i, _p, _p, _p, _p]),
    "b200grpo_token_loss_workspace": (_sz, [_i, _i]),
    "b200grpo_token_loss": (_i, [_p] * 12 + [_i, _i, C.POINTER(LossCfg)] + [_p] * 6 + [_sz, _p]),
    "b200grpo_segment_bounds_workspace": (_sz, [_i]),
    "b200grpo_segment_bounds": (_i, [_p, _i, _p, _p, _p, _sz, _p]),
    "b200grpo_segment_bounds_static": (_i, [_p, _i, _p, _p, _p]),   # scratch inside the library; allocates nothing
    "b200grpo_lmhead_bwd_workspace": (_sz, [_i, _i, _i, _i]),
    "b200grpo_lmhead_bwd": (_i, [_p, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i64, _f, _p, _p, _i64, _p, _i64, _i, _i,
                                 _p, _sz, _p]),
    "b200grpo_lmhead_bwd_staged": (_i, [_p, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i64, _f, _p, _p, _i64, _p, _i64, _i, _i,
                                        _i, _p, _sz, _p]),
    "b200grpo_group_advantage_workspace": (_sz, [_i, _i]),
    "b200grpo_group_advantage": (_i, [_p] * 8 + [_i, _i, _i] + [_p] * 6 + [_sz, _p]),
    "b200grpo_group_advantage_rows": (_i, [_p] * 5 + [_i, _i, _i] + [_p] * 3 + [_sz, _p]),
    "b200grpo_bucket_workspace": (_sz, [_i]),
    "b200grpo_bucket_numel": (_i64, [_p, _i]),
    "b200grpo_bucket_pack": (_i, [_p, _p, _p, _i, _p, _i64, _p, _sz, _p]),
    "b200grpo_bucket_pack_multicast": (_i, [_p, _p, _p, _i, _p, _i64, _p, _sz, _p]),
    "b200grpo_bucket_scatter_workspace": (_sz, [_i, _i]),
    "b200grpo_bucket_pack_scatter": (_i, [_p, _p, _p, _i, _p, _i, _i64, _i64, _p, _sz, _p]),
    "b200grpo_bucket_relay_multicast": (_i, [_p, _p, _i64, _i, _i, _i64, _i, _p]),
    "b200grpo_bucket_relay_p2p": (_i, [_p, _p, _i64, _i, _i, _i64, _i, _p]),
    "b200grpo_bucket_unpack": (_i, [_p, _p, _p, _p, _i, _p, _sz, _p]),
    "b200grpo_rl_step_host_workspace": (_sz, [_i, _i, _i]),
    "b200grpo_rl_step_host_grad_hidden": (_p, [_p, _i, _i, _i]),
    "b200grpo_rl_step_host": (_i, [C.POINTER(HostBatch), _i, _i, _i, _p, _i64, _f, C.POINTER(LossCfg), _p, _p, _i, _i, _p, _p,
                                   _sz, _p]),
    "b200grpo_expand_packed": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, C.POINTER(PackedOut), _p]),
    "b200grpo_metrics_reset": (_i, [_p, _i, _p, _p]),
    "b200grpo_metrics_accumulate": (_i, [_p, _p, _p, _i, _d, _d, _d, _d, _i, _p, _p]),
    "b200grpo_final_norm_bwd_workspace": (_sz, [_i, _i]),
    "b200grpo_final_norm_fwd": (_i, [_p, _i64, _p, _f, _p, _p, _i, _i, _p, _i64, _p, _p, _p]),
    "b200grpo_final_norm_bwd": (_i, [_p, _i64, _p, _p, _p, _i64, _p, _p, _i, _i, _p, _i64, _p, _p, _p, _p, _sz, _p]),
    "b200grpo_profile_enable": (_i, [_i]),
    "b200grpo_profile_read": (_i, [_p, _p, _i]),
    "b200grpo_gemm_bf16": (_i, [_p, _i64, _i, _p, _i64, _i, _p, _i64, _i, _i, _i, _i, _i, _i, _i, _p, _sz, _p]),
}

_lib = None
_lock = threading.Lock()


def lib() -> C.CDLL:
    """Load (once) and return the library; raise loudly if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise B200GrpoError(
                    f"{LIB_PATH} is not built. pipelinerl_swe_b200 has no CPU or PyTorch fallback: build the CUDA "
                    "library with `make -C pipelinerl_swe_b200/csrc` (or __graft_entry__.build()).")
            handle = C.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(handle, name)
                fn.restype, fn.argtypes = res, args
            # experiments: B200GRPO_TUNABLES="name=value,name=value" applies library tunables at load time, so that a whole
            # test or bench run can be repeated under a non-default kernel schedule
            for kv in filter(None, os.environ.get("B200GRPO_TUNABLES", "").split(",")):
                name, _, val = kv.partition("=")
                if handle.b200grpo_set_tunable(name.strip().encode(), int(val)) < 0:   # returns the previous value (>= 0) or an error code
                    raise B200GrpoError(f"B200GRPO_TUNABLES: unknown tunable {name!r}")
            _lib = handle
    return _lib


def check(code: int) -> None:
    if code != 0:
        msg = lib().b200grpo_last_error().decode(errors="replace")
        raise B200GrpoError(f"libb200grpo error {code}: {msg}")


PROF_TAGS = {1: "fwd_gemm_stats", 2: "fwd_merge", 3: "finalize", 4: "token_loss", 5: "bwd_recompute_dlogits", 6: "bwd_dH_gemm",
             7: "bwd_dW_gemm", 8: "gemm", 9: "group_advantage", 10: "bucket", 11: "segments", 12: "fused_logits_gemm",
             13: "dlogits_pass", 14: "splitk_reduce", 15: "ingest_expand", 16: "final_norm", 17: "mc_allreduce"}


def profile_enable(capacity: int) -> None:
    check(lib().b200grpo_profile_enable(int(capacity)))


def profile_read(capacity: int = 4096) -> list[tuple[str, float]]:
    tags = (C.c_int32 * capacity)()
    ms = (C.c_float * capacity)()
    n = lib().b200grpo_profile_read(tags, ms, capacity)
    if n < 0:
        check(n)
    return [(PROF_TAGS.get(tags[i], str(tags[i])), float(ms[i])) for i in range(n)]


def launch_count() -> int:
    return int(lib().b200grpo_launch_count())


def set_tunable(name: str, value: int) -> int:
    return lib().b200grpo_set_tunable(name.encode(), int(value))
