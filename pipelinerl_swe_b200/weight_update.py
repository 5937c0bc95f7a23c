"""Bucketed trainer -> actor weight broadcast (the receive_weight_update path).

Reference behaviour (pipelinerl/finetune_loop.py:236-248, vllm1.py:81-94): for every
parameter, the sender casts to bf16 and broadcasts it from rank 0 of the actor-update
group; every receiver allocates a fresh buffer, joins the broadcast and calls
``load_weights([(name, buffer)])`` -- 339 launches + allocations for Qwen2.5-7B.

Here the parameter list (same ``ParameterInfo`` schema, same order) is cut into buckets of
``bucket_bytes``; K6 packs a bucket (fused fp32->bf16 cast) into one of two persistent
staging buffers, ONE ``dist.broadcast`` moves it over NVLink/NVSwitch on a side stream
while the next bucket is being packed, and receivers hand each parameter to
``load_weights`` as a zero-copy view of the received bucket.  The wire content per
parameter is byte-identical to the reference's ``param.data.bfloat16()``.
"""
from __future__ import annotations

import time
from typing import Callable, Iterable, Literal, Sequence

import torch
import torch.distributed as dist
from pydantic import BaseModel

from . import ops


class ParameterInfo(BaseModel):  # finetune_loop.py:122-125
    name: str
    shape: list[int]
    dtype: str


class WeightUpdateRequest(BaseModel):  # finetune_loop.py:128-132
    kind: Literal["weight_update_request"] = "weight_update_request"
    version: int
    parameters_info: list[ParameterInfo]
    timestamp: float = time.time()


def _numel(shape: Sequence[int]) -> int:
    n = 1
    for s in shape:
        n *= int(s)
    return n


def plan_buckets(parameters_info: Sequence[ParameterInfo], bucket_bytes: int) -> list[list[int]]:
    """Greedy in-order split; a parameter larger than the bucket gets a bucket of its own.

    Sender and receivers derive the same plan from the same WeightUpdateRequest, so no
    extra metadata crosses the wire.  Sizes follow K6's 16-byte alignment rule.
    """
    cap = max(int(bucket_bytes) // 2, 8)
    buckets: list[list[int]] = []
    cur: list[int] = []
    used = 0
    for i, info in enumerate(parameters_info):
        n = (_numel(info.shape) + 7) // 8 * 8
        if cur and used + n > cap:
            buckets.append(cur)
            cur, used = [], 0
        cur.append(i)
        used += n
    if cur:
        buckets.append(cur)
    return buckets


def _bucket_elems(parameters_info, idxs) -> int:
    return sum((_numel(parameters_info[i].shape) + 7) // 8 * 8 for i in idxs)


class _Staging:
    """Two persistent bf16 staging buffers + a side stream (double buffering)."""

    def __init__(self, device, elems: int):
        self.bufs = [torch.empty(elems, dtype=torch.bfloat16, device=device) for _ in range(2)]
        self.free = [torch.cuda.Event() for _ in range(2)]
        self.comm = torch.cuda.Stream(device=device)
        for ev in self.free:
            ev.record()


def make_request(named_parameters: Iterable[tuple[str, torch.Tensor]], version: int) -> WeightUpdateRequest:
    """ParameterInfo list exactly as the reference builds it (dtype string 'torch.bfloat16')."""
    infos = [ParameterInfo(name=n, shape=list(p.shape), dtype=str(torch.bfloat16)) for n, p in named_parameters]
    return WeightUpdateRequest(version=version, parameters_info=infos)


def send_weight_update(named_parameters: Sequence[tuple[str, torch.Tensor]], group, bucket_bytes: int = 256 << 20,
                       src: int = 0, staging: _Staging | None = None) -> _Staging:
    """Sender side (finetune_loop.py:236-248): pack -> broadcast per bucket, then barrier."""
    params = [p.detach() for _, p in named_parameters]
    params = [p if p.is_contiguous() else p.contiguous() for p in params]
    infos = [ParameterInfo(name=n, shape=list(p.shape), dtype=str(torch.bfloat16)) for n, p in named_parameters]
    buckets = plan_buckets(infos, bucket_bytes)
    need = max(_bucket_elems(infos, b) for b in buckets)
    dev = params[0].device
    if staging is None or staging.bufs[0].numel() < need:
        staging = _Staging(dev, need)
    main = torch.cuda.current_stream(dev)
    for k, idxs in enumerate(buckets):
        slot = k & 1
        n = _bucket_elems(infos, idxs)
        buf = staging.bufs[slot][:n]
        main.wait_event(staging.free[slot])           # previous broadcast out of this slot finished
        ops.bucket_pack([params[i] for i in idxs], buf)
        packed = torch.cuda.Event()
        packed.record(main)
        with torch.cuda.stream(staging.comm):
            staging.comm.wait_event(packed)
            dist.broadcast(buf, src=src, group=group)
            staging.free[slot].record(staging.comm)
    main.wait_stream(staging.comm)
    dist.barrier(group)
    return staging


def receive_weight_update(request: WeightUpdateRequest, group, load_weights: Callable[[list[tuple[str, torch.Tensor]]], object],
                          device, model_dtype: torch.dtype = torch.bfloat16, bucket_bytes: int = 256 << 20, src: int = 0,
                          staging: _Staging | None = None) -> _Staging:
    """Receiver side (vllm1.py:81-94): same checks, bucketed broadcasts, zero-copy views into the bucket."""
    infos = request.parameters_info
    for info in infos:
        assert info.dtype == str(model_dtype), f"mismatch dtype: src {info.dtype}, dst {model_dtype}"
    buckets = plan_buckets(infos, bucket_bytes)
    need = max(_bucket_elems(infos, b) for b in buckets)
    if staging is None or staging.bufs[0].numel() < need:
        staging = _Staging(device, need)
    main = torch.cuda.current_stream(device)
    for k, idxs in enumerate(buckets):
        slot = k & 1
        n = _bucket_elems(infos, idxs)
        buf = staging.bufs[slot][:n]
        with torch.cuda.stream(staging.comm):
            staging.comm.wait_event(staging.free[slot])   # loads out of this slot finished
            dist.broadcast(buf, src=src, group=group)
            arrived = torch.cuda.Event()
            arrived.record(staging.comm)
        main.wait_event(arrived)
        off = 0
        for i in idxs:
            numel = _numel(infos[i].shape)
            view = buf[off:off + numel].view(tuple(infos[i].shape))
            loaded = load_weights([(infos[i].name, view)])
            if loaded is not None and len(loaded) != 1:
                raise ValueError(f"model {infos[i].name} not found in model state dict")
            off += (numel + 7) // 8 * 8
        staging.free[slot].record(main)
    dist.barrier(group)
    return staging


def reference_send(named_parameters, group, src: int = 0) -> None:
    """The reference's per-parameter loop, restated for A/B timing (finetune_loop.py:245-248)."""
    for _, p in named_parameters:
        dist.broadcast(p.data.bfloat16(), src=src, group=group)
    dist.barrier(group)


def reference_receive(request: WeightUpdateRequest, group, load_weights, device, src: int = 0) -> None:
    """vllm1.py:81-94 restated: fresh buffer + broadcast + load per parameter."""
    for info in request.parameters_info:
        buffer = torch.empty(tuple(info.shape), dtype=torch.bfloat16, device=device)
        dist.broadcast(buffer, src=src, group=group)
        load_weights([(info.name, buffer)])
    dist.barrier(group)
