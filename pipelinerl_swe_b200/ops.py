"""Tensor-level wrappers over the C ABI: torch owns memory and streams, the library computes.

Each wrapper validates dtype / contiguity, sizes the workspace with the library's own
``*_workspace`` query, and passes raw device pointers plus the current CUDA stream.
No wrapper computes anything in torch.
"""
from __future__ import annotations

import ctypes as C
from typing import Sequence

import torch

from . import _abi
from ._abi import LossCfg, M_COUNT, check, lib

_ws_cache: dict[tuple[int, str], torch.Tensor] = {}


def _first_cuda_tensor(args):
    for a in args:
        if isinstance(a, torch.Tensor):
            if a.is_cuda:
                return a
        elif isinstance(a, (list, tuple)):
            t = _first_cuda_tensor(a)
            if t is not None:
                return t
        elif isinstance(a, dict):
            t = _first_cuda_tensor(a.values())
            if t is not None:
                return t
    return None


def _on_tensor_device(fn):
    """Run ``fn`` with the device of its first CUDA tensor argument current: the library launches on the CURRENT device and
    stream, so a caller that holds tensors on cuda:1 while cuda:0 is current (one process driving several GPUs, a model moved
    without ``torch.cuda.set_device``) is served like torch serves it.  One integer comparison when the device already is current."""
    import functools

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        t = _first_cuda_tensor(args)
        if t is None:
            t = _first_cuda_tensor(kwargs.values())
        if t is None or t.device.index == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(t.device):
            return fn(*args, **kwargs)
    return wrapper


def _workspace(device: torch.device, nbytes: int, tag: str = "main") -> torch.Tensor:
    """Grow-only per-device scratch; calls are stream-ordered so one buffer per tag suffices."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), tag)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _req(t: torch.Tensor, dtype: torch.dtype, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise _abi.B200GrpoError(f"{name} must be a CUDA tensor (pipelinerl_swe_b200 has no CPU path)")
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    return t


def lmhead_saved_buffer(M: int, H: int, V: int, device) -> torch.Tensor:
    """Scratch the forward fills for a recompute-free backward (bf16 logits relative to their tile max)."""
    n = lib().b200grpo_lmhead_saved_bytes(M, H, V)
    return torch.empty(n, dtype=torch.uint8, device=device)


@_on_tensor_device
def lmhead_fwd_stats(hidden: torch.Tensor, weight: torch.Tensor, targets: torch.Tensor, inv_temperature: float,
                     v_offset: int = 0, saved: torch.Tensor | None = None, stats_mc: int | None = None) -> torch.Tensor | None:
    """K1: per-row (max, sumexp, sum e*z, z_target) over this rank's vocab rows -> f32 [M,4].

    ``stats_mc``: NVSwitch multicast address of this rank's [M,4] slot in a symmetric buffer -- the merge kernel stores
    there with multimem.st (the vocab-parallel all-gather is the kernel's own store); nothing is returned."""
    _req(hidden, torch.bfloat16, "hidden"); _req(weight, torch.bfloat16, "weight"); _req(targets, torch.int64, "targets")
    M, H = hidden.shape
    V = weight.shape[0]
    assert weight.shape[1] == H and targets.numel() == M
    n = lib().b200grpo_lmhead_fwd_workspace(M, H, V)
    ws = _workspace(hidden.device, n)
    if stats_mc is not None:
        check(lib().b200grpo_lmhead_fwd_multicast(hidden.data_ptr(), H, weight.data_ptr(), H, targets.data_ptr(), M, H, V, v_offset,
                                                  float(inv_temperature), C.c_void_p(stats_mc), _ptr(saved), ws.data_ptr(), ws.numel(),
                                                  _stream()))
        return None
    stats = torch.empty(M, 4, dtype=torch.float32, device=hidden.device)
    check(lib().b200grpo_lmhead_fwd(hidden.data_ptr(), H, weight.data_ptr(), H, targets.data_ptr(), M, H, V, v_offset,
                                    float(inv_temperature), stats.data_ptr(), _ptr(saved), ws.data_ptr(), ws.numel(), _stream()))
    return stats


@_on_tensor_device
def lmhead_finalize(stats: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """stats f32 [R,M,4] (or [M,4]) -> (logprob, entropy, lse) f32 [M]."""
    _req(stats, torch.float32, "stats")
    if stats.dim() == 2:
        stats = stats[None]
    R, M, _ = stats.shape
    out = torch.empty(3, M, dtype=torch.float32, device=stats.device)
    check(lib().b200grpo_lmhead_finalize(stats.data_ptr(), R, M, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), _stream()))
    return out[0], out[1], out[2]


@_on_tensor_device
def lmhead_bwd(hidden, weight, targets, lse, entropy, coef_lp, coef_ent, inv_temperature, v_offset=0, grad_scale=None,
               need_grad_hidden=True, need_grad_weight=True, grad_weight_out: torch.Tensor | None = None,
               accumulate: bool = False, saved: torch.Tensor | None = None, stages: int | None = None,
               grad_hidden_out: torch.Tensor | None = None):
    """K2: (grad_hidden bf16 [M,H] | None, grad_weight [V,H] | None).  ``saved`` (from the forward) is consumed.

    ``coef_ent=None`` declares that no gradient flows through the entropy (``entropy_bonus == 0``): with ``saved`` the
    backward then runs its two GEMMs straight on the forward's saved exponentials (no [tokens x vocab] dlogits pass).

    ``stages`` (``_abi.BWD_*`` bits) splits the backward: a first call with ``BWD_DLOGITS | BWD_DH`` and a second with
    ``BWD_DW`` (same ``saved``, nothing re-formed) let the caller put a collective on grad_hidden in between."""
    _req(hidden, torch.bfloat16, "hidden"); _req(weight, torch.bfloat16, "weight"); _req(targets, torch.int64, "targets")
    if coef_ent is None and saved is None:
        coef_ent = torch.zeros_like(coef_lp)     # the recompute GEMM's epilogue takes the coefficient explicitly
    for name, t in (("lse", lse), ("entropy", entropy), ("coef_lp", coef_lp), ("coef_ent", coef_ent)):
        if t is not None:
            _req(t, torch.float32, name)
    M, H = hidden.shape
    V = weight.shape[0]
    if stages is not None:
        need_grad_hidden = need_grad_hidden and bool(stages & _abi.BWD_DH)
        need_grad_weight = need_grad_weight and bool(stages & _abi.BWD_DW)
    gh = None
    if need_grad_hidden:
        gh = grad_hidden_out if grad_hidden_out is not None else torch.empty(M, H, dtype=torch.bfloat16, device=hidden.device)
    gw = None
    if need_grad_weight:
        gw = grad_weight_out if grad_weight_out is not None else torch.empty(V, H, dtype=torch.bfloat16, device=hidden.device)
        assert gw.shape == (V, H) and gw.is_contiguous() and gw.dtype in (torch.bfloat16, torch.float32)
    n = lib().b200grpo_lmhead_bwd_workspace(M, H, V, int(saved is not None))
    ws = _workspace(hidden.device, n, "dlogits")
    args = (hidden.data_ptr(), H, weight.data_ptr(), H, targets.data_ptr(), lse.data_ptr(), entropy.data_ptr(), coef_lp.data_ptr(),
            _ptr(coef_ent), _ptr(grad_scale), M, H, V, v_offset, float(inv_temperature), _ptr(saved), _ptr(gh), H, _ptr(gw), H,
            int(gw is not None and gw.dtype == torch.float32), int(accumulate))
    if stages is None:
        check(lib().b200grpo_lmhead_bwd(*args, ws.data_ptr(), ws.numel(), _stream()))
    else:
        check(lib().b200grpo_lmhead_bwd_staged(*args, int(stages), ws.data_ptr(), ws.numel(), _stream()))
    return gh, gw


@_on_tensor_device
def token_loss(logprob, entropy, batch: dict, cfg: LossCfg, values=None):
    """K3 -> (loss f32[1], metrics f64[M_COUNT], coef_lp, coef_ent, coef_value|None)."""
    labels = _req(batch["labels"], torch.int64, "labels")
    rows, L = labels.shape
    dev = labels.device
    f32 = lambda k: _req(batch[k], torch.float32, k)  # noqa: E731
    pos = batch.get("position_ids")
    if pos is not None:
        _req(pos, torch.int64, "position_ids")
    _req(logprob, torch.float32, "logprob"); _req(entropy, torch.float32, "entropy")
    assert logprob.numel() == rows * (L - 1) and entropy.numel() == rows * (L - 1)
    metrics = torch.empty(M_COUNT, dtype=torch.float64, device=dev)
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    coef = torch.empty(2, rows, max(L - 1, 0), dtype=torch.float32, device=dev)
    coef_v = torch.empty(rows, L, dtype=torch.float32, device=dev) if values is not None else None
    if values is not None:
        _req(values, torch.float32, "values")
    n = lib().b200grpo_token_loss_workspace(rows, L)
    ws = _workspace(dev, n)
    check(lib().b200grpo_token_loss(
        logprob.data_ptr(), entropy.data_ptr(), labels.data_ptr(), _ptr(pos), f32("rewards").data_ptr(),
        f32("advantages").data_ptr(), f32("ref_logprobs").data_ptr(), f32("old_logprobs").data_ptr(),
        f32("group_tokens").data_ptr(), f32("num_labels").data_ptr(), f32("overflow").data_ptr(), _ptr(values), rows, L,
        C.byref(cfg), metrics.data_ptr(), loss.data_ptr(), coef[0].data_ptr(), coef[1].data_ptr(), _ptr(coef_v),
        ws.data_ptr(), ws.numel(), _stream()))
    return loss, metrics, coef[0], coef[1], coef_v


@_on_tensor_device
def segment_bounds(position_ids: torch.Tensor) -> torch.Tensor:
    """int32 boundaries [S+1] of a packed row, as rl/__init__.py:159-170 (one D2H sync for S)."""
    pos = _req(position_ids.reshape(-1), torch.int64, "position_ids")
    T = pos.numel()
    bounds = torch.empty(T + 1, dtype=torch.int32, device=pos.device)
    n = torch.empty(1, dtype=torch.int32, device=pos.device)
    ws = _workspace(pos.device, lib().b200grpo_segment_bounds_workspace(T))
    check(lib().b200grpo_segment_bounds(pos.data_ptr(), T, bounds.data_ptr(), n.data_ptr(), ws.data_ptr(), ws.numel(), _stream()))
    return bounds[: int(n.item())]


@_on_tensor_device
def gemm_bf16(A: torch.Tensor, B: torch.Tensor, a_mn_major=False, b_mn_major=False, out_dtype=torch.float32,
              out: torch.Tensor | None = None, accumulate=False, group_m: int = 0, k_splits: int = 1) -> torch.Tensor:
    """C[M,N] = A.B^T on the tcgen05 mainloop (test surface).  MN-major operands are given as [K, M|N]."""
    _req(A, torch.bfloat16, "A"); _req(B, torch.bfloat16, "B")
    (K, M) = A.shape if a_mn_major else A.shape[::-1]
    (K2, N) = B.shape if b_mn_major else B.shape[::-1]
    assert K == K2
    if out is None:
        out = torch.empty(M, N, dtype=out_dtype, device=A.device)
    # k_splits == 0: tail split of the last partial wave, at most one f32 tile per cluster (<= 74 x 512 KB)
    ws = _workspace(A.device, k_splits * M * N * 4 if k_splits > 1 else (40 << 20) if k_splits == 0 else 256, "slab")
    check(lib().b200grpo_gemm_bf16(A.data_ptr(), A.stride(0), int(a_mn_major), B.data_ptr(), B.stride(0), int(b_mn_major),
                                   out.data_ptr(), out.stride(0), int(out.dtype == torch.float32), int(accumulate), M, N, K,
                                   int(group_m), int(k_splits), ws.data_ptr(), ws.numel(), _stream()))
    return out


@_on_tensor_device
def group_order(group_key: torch.Tensor, n_groups: int) -> tuple[torch.Tensor, torch.Tensor]:
    """(row_order i32[n_rows], group_starts i32[n_groups+1]): rows stably sorted by dense group key -- the CSR K4 walks."""
    order = torch.argsort(group_key, stable=True).to(torch.int32)
    starts = torch.zeros(n_groups + 1, dtype=torch.int32, device=group_key.device)
    if group_key.numel():
        starts[1:] = torch.cumsum(torch.bincount(group_key, minlength=n_groups), 0).to(torch.int32)
    return order, starts


@_on_tensor_device
def group_advantage(row_reward: torch.Tensor, group_key: torch.Tensor, row_offsets: torch.Tensor, n_labels: torch.Tensor,
                    has_eos: torch.Tensor, n_groups: int, divide_by_std: bool, token_rewards: torch.Tensor | None = None,
                    row_order: torch.Tensor | None = None, group_starts: torch.Tensor | None = None, want_f64: bool = False):
    """K4 -> f32 [4, total_tokens]: advantages, group_tokens, num_labels, overflow (+ f64 [total] advantages if ``want_f64``)."""
    _req(row_reward, torch.float64, "row_reward"); _req(group_key, torch.int32, "group_key")
    _req(row_offsets, torch.int64, "row_offsets"); _req(n_labels, torch.int32, "n_labels"); _req(has_eos, torch.uint8, "has_eos")
    if token_rewards is not None:
        _req(token_rewards, torch.float64, "token_rewards")
    if row_order is None:
        row_order, group_starts = group_order(group_key, n_groups)
    _req(row_order, torch.int32, "row_order"); _req(group_starts, torch.int32, "group_starts")
    n_rows = row_reward.numel()
    total = int(row_offsets[-1].item()) if n_rows else 0
    pitch = (total + 3) // 4 * 4  # 16-byte aligned columns -> vector stores in the expand kernel
    out = torch.empty(4, pitch, dtype=torch.float32, device=row_reward.device)[:, :total]
    adv64 = torch.empty(total, dtype=torch.float64, device=row_reward.device) if want_f64 else None
    n = lib().b200grpo_group_advantage_workspace(n_rows, n_groups)
    ws = _workspace(row_reward.device, n)
    check(lib().b200grpo_group_advantage(row_reward.data_ptr(), group_key.data_ptr(), row_order.data_ptr(), group_starts.data_ptr(),
                                         row_offsets.data_ptr(), n_labels.data_ptr(), has_eos.data_ptr(), _ptr(token_rewards), n_rows,
                                         n_groups, int(divide_by_std), out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                                         out[3].data_ptr(), _ptr(adv64), ws.data_ptr(), ws.numel(), _stream()))
    return (out, adv64) if want_f64 else out


@_on_tensor_device
def group_advantage_rows(row_reward: torch.Tensor, group_key: torch.Tensor, row_offsets: torch.Tensor, n_groups: int,
                         divide_by_std: bool, row_order: torch.Tensor | None = None,
                         group_starts: torch.Tensor | None = None) -> torch.Tensor:
    """K4, per-rollout form -> f64 [2, n_rows]: advantage and mean group length of every rollout (bit-identical to pandas)."""
    _req(row_reward, torch.float64, "row_reward"); _req(group_key, torch.int32, "group_key")
    _req(row_offsets, torch.int64, "row_offsets")
    if row_order is None:
        row_order, group_starts = group_order(group_key, n_groups)
    _req(row_order, torch.int32, "row_order"); _req(group_starts, torch.int32, "group_starts")
    n_rows = row_reward.numel()
    out = torch.empty(2, n_rows, dtype=torch.float64, device=row_reward.device)
    ws = _workspace(row_reward.device, lib().b200grpo_group_advantage_workspace(n_rows, n_groups))
    check(lib().b200grpo_group_advantage_rows(row_reward.data_ptr(), group_key.data_ptr(), row_order.data_ptr(), group_starts.data_ptr(),
                                              row_offsets.data_ptr(), n_rows, n_groups, int(divide_by_std), out[0].data_ptr(),
                                              out[1].data_ptr(), ws.data_ptr(), ws.numel(), _stream()))
    return out


def _tensor_table(tensors: Sequence[torch.Tensor]):
    n = len(tensors)
    ptrs = (C.c_void_p * n)(*[t.data_ptr() for t in tensors])
    numel = (C.c_int64 * n)(*[t.numel() for t in tensors])
    is_f32 = (C.c_uint8 * n)(*[int(t.dtype == torch.float32) for t in tensors])
    for t in tensors:
        if t.dtype not in (torch.bfloat16, torch.float32) or not t.is_contiguous() or not t.is_cuda:
            raise TypeError("bucket tensors must be contiguous CUDA bf16 or f32")
    return ptrs, numel, is_f32


def bucket_numel(tensors: Sequence[torch.Tensor]) -> int:
    n = len(tensors)
    numel = (C.c_int64 * n)(*[t.numel() for t in tensors])
    return int(lib().b200grpo_bucket_numel(numel, n))


@_on_tensor_device
def bucket_pack(tensors: Sequence[torch.Tensor], bucket: torch.Tensor) -> None:
    """K6: gather parameters (bf16, or f32 cast on the fly) into one contiguous bf16 bucket."""
    _req(bucket, torch.bfloat16, "bucket")
    ptrs, numel, is_f32 = _tensor_table(tensors)
    n = len(tensors)
    ws = _workspace(bucket.device, lib().b200grpo_bucket_workspace(n), "bucket")
    check(lib().b200grpo_bucket_pack(ptrs, numel, is_f32, n, bucket.data_ptr(), bucket.numel(), ws.data_ptr(), ws.numel(), _stream()))


@_on_tensor_device
def bucket_pack_multicast(tensors: Sequence[torch.Tensor], mc_ptr: int, capacity: int) -> None:
    """K6 + broadcast in one pass: pack into the NVSwitch multicast mapping ``mc_ptr`` (bf16 elements, ``capacity`` of them)."""
    ptrs, numel, is_f32 = _tensor_table(tensors)
    n = len(tensors)
    ws = _workspace(tensors[0].device, lib().b200grpo_bucket_workspace(n), "bucket")
    check(lib().b200grpo_bucket_pack_multicast(ptrs, numel, is_f32, n, C.c_void_p(mc_ptr), int(capacity), ws.data_ptr(), ws.numel(), _stream()))


@_on_tensor_device
def bucket_pack_scatter(tensors: Sequence[torch.Tensor], peer_ptrs: Sequence[int], stripe_elems: int, capacity: int) -> None:
    """K6 + scatter in one pass: stripe j of the packed bucket is stored into ``peer_ptrs[j % len(peer_ptrs)]`` (device
    pointers to the receivers' copies of a symmetric slot, P2P-mapped) -- the first half of the scatter + relay broadcast."""
    ptrs, numel, is_f32 = _tensor_table(tensors)
    n, q = len(tensors), len(peer_ptrs)
    peers = (C.c_void_p * q)(*[int(p) for p in peer_ptrs])
    ws = _workspace(tensors[0].device, lib().b200grpo_bucket_scatter_workspace(n, q), "bucket")
    check(lib().b200grpo_bucket_pack_scatter(ptrs, numel, is_f32, n, peers, q, int(stripe_elems), int(capacity), ws.data_ptr(),
                                             ws.numel(), _stream()))


def bucket_relay_multicast(local_ptr: int, mc_ptr: int, numel: int, peer_index: int, n_peers: int, stripe_elems: int,
                           ctas: int = 0) -> None:
    """Second half of scatter + relay: re-emit the stripes this receiver owns from its local copy through the NVSwitch
    multicast address of the slot (multimem.st), completing every other rank's copy."""
    check(lib().b200grpo_bucket_relay_multicast(C.c_void_p(local_ptr), C.c_void_p(mc_ptr), int(numel), int(peer_index), int(n_peers),
                                                int(stripe_elems), int(ctas), _stream()))


def bucket_relay_p2p(local_ptr: int, peer_ptrs: Sequence[int], numel: int, peer_index: int, stripe_elems: int, ctas: int = 0) -> None:
    """The relay over plain P2P stores: the owned stripes are written into every other receiver's copy of the slot."""
    q = len(peer_ptrs)
    peers = (C.c_void_p * q)(*[int(p) for p in peer_ptrs])
    check(lib().b200grpo_bucket_relay_p2p(C.c_void_p(local_ptr), peers, int(numel), int(peer_index), q, int(stripe_elems), int(ctas), _stream()))


@_on_tensor_device
def bucket_unpack(bucket: torch.Tensor, tensors: Sequence[torch.Tensor]) -> None:
    """K6 inverse: scatter a received bucket into the destination parameters."""
    _req(bucket, torch.bfloat16, "bucket")
    ptrs, numel, is_f32 = _tensor_table(tensors)
    n = len(tensors)
    ws = _workspace(bucket.device, lib().b200grpo_bucket_workspace(n), "bucket")
    check(lib().b200grpo_bucket_unpack(bucket.data_ptr(), ptrs, numel, is_f32, n, ws.data_ptr(), ws.numel(), _stream()))


@_on_tensor_device
def final_norm_fwd(x: torch.Tensor, gamma: torch.Tensor, eps: float, value_w: torch.Tensor | None = None,
                   value_b: torch.Tensor | None = None) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor | None]:
    """Final RMSNorm (+ value head) in one sweep: x bf16 [T,H] -> (y bf16 [T,H], rstd f32 [T], value f32 [T] | None)."""
    _req(x, torch.bfloat16, "x"); _req(gamma, torch.bfloat16, "gamma")
    T, H = x.shape
    assert gamma.numel() == H
    if value_w is not None:
        _req(value_w, torch.bfloat16, "value_w")
        assert value_w.numel() == H
        if value_b is not None:
            _req(value_b, torch.bfloat16, "value_b")
    y = torch.empty_like(x)
    rstd = torch.empty(T, dtype=torch.float32, device=x.device)
    value = torch.empty(T, dtype=torch.float32, device=x.device) if value_w is not None else None
    check(lib().b200grpo_final_norm_fwd(x.data_ptr(), H, gamma.data_ptr(), float(eps), _ptr(value_w), _ptr(value_b), T, H,
                                        y.data_ptr(), H, rstd.data_ptr(), _ptr(value), _stream()))
    return y, rstd, value


@_on_tensor_device
def final_norm_bwd(x: torch.Tensor, gamma: torch.Tensor, rstd: torch.Tensor, dy: torch.Tensor,
                   dvalue: torch.Tensor | None = None, value_w: torch.Tensor | None = None):
    """-> (dx bf16 [T,H], dgamma f32 [H], dvalue_w f32 [H] | None, dvalue_b f32 [1] | None)."""
    _req(x, torch.bfloat16, "x"); _req(gamma, torch.bfloat16, "gamma"); _req(rstd, torch.float32, "rstd"); _req(dy, torch.bfloat16, "dy")
    T, H = x.shape
    assert dy.shape == x.shape and rstd.numel() == T
    dx = torch.empty_like(x)
    dgamma = torch.empty(H, dtype=torch.float32, device=x.device)
    dvw = dvb = None
    if dvalue is not None:
        _req(dvalue, torch.float32, "dvalue"); _req(value_w, torch.bfloat16, "value_w")
        dvw = torch.empty(H, dtype=torch.float32, device=x.device)
        dvb = torch.empty(1, dtype=torch.float32, device=x.device)
    n = lib().b200grpo_final_norm_bwd_workspace(T, H)
    ws = _workspace(x.device, n, "final_norm")
    check(lib().b200grpo_final_norm_bwd(x.data_ptr(), H, gamma.data_ptr(), rstd.data_ptr(), dy.data_ptr(), H, _ptr(dvalue),
                                        _ptr(value_w), T, H, dx.data_ptr(), H, dgamma.data_ptr(), _ptr(dvw), _ptr(dvb),
                                        ws.data_ptr(), ws.numel(), _stream()))
    return dx, dgamma, dvw, dvb


def mc_allreduce_bf16(mc_ptr: int, numel: int, rank: int, world: int, ctas: int = 16) -> None:
    """In place sum over ranks of a symmetric bf16 tensor through its NVSwitch multicast address (two-shot, in-switch reduce)."""
    check(lib().b200grpo_mc_allreduce_bf16(C.c_void_p(mc_ptr), int(numel), int(rank), int(world), int(ctas), _stream()))
