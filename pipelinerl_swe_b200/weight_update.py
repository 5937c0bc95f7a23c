"""Bucketed trainer -> actor weight broadcast (the receive_weight_update path).

Reference behaviour (pipelinerl/finetune_loop.py:236-248, vllm1.py:81-94): for every
parameter, the sender casts to bf16 and broadcasts it from rank 0 of the actor-update
group; every receiver allocates a fresh buffer, joins the broadcast and calls
``load_weights([(name, buffer)])`` -- 339 launches + allocations for Qwen2.5-7B.

Here the parameter list (same ``ParameterInfo`` schema) is split by ``plan_transfers``: large
tensors are broadcast on their own, straight out of the bf16 parameter storage; the many
small ones are packed by K6 (fused fp32->bf16 cast) into buckets of ``bucket_bytes`` held in
two persistent staging buffers.  Broadcasts run on a side stream over NVLink/NVSwitch while
the next bucket is being packed, and receivers hand each parameter to ``load_weights`` as a
zero-copy view of the received buffer (double-buffered, so load k overlaps broadcast k+1).  The wire content per
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


def plan_transfers(parameters_info: Sequence[ParameterInfo], bucket_bytes: int, direct_bytes: int = 8 << 20):
    """Wire plan both sides derive from the same request: [("direct", i) | ("bucket", [i, ...]), ...].

    Tensors of at least ``direct_bytes`` are broadcast on their own (a bf16 parameter goes straight from its
    storage, zero-copy); the many small ones (biases, norms, GQA k/v projections: 197 of Qwen2.5-7B's 339
    tensors but 1.4 % of its bytes) are gathered by K6 into buckets of ``bucket_bytes`` so that their
    per-broadcast launch latency is paid once per bucket.  Order inside each class follows the request.
    """
    small = [i for i, info in enumerate(parameters_info) if 2 * _numel(info.shape) < direct_bytes]
    big = [i for i, info in enumerate(parameters_info) if 2 * _numel(info.shape) >= direct_bytes]
    plan: list[tuple[str, object]] = []
    if small:
        sub = [parameters_info[i] for i in small]
        plan += [("bucket", [small[j] for j in b]) for b in plan_buckets(sub, bucket_bytes)]
    plan += [("direct", i) for i in big]
    return plan


def _staging_need(infos, plan) -> int:
    need = 8
    for kind, item in plan:
        need = max(need, _bucket_elems(infos, item) if kind == "bucket" else (_numel(infos[item].shape) + 7) // 8 * 8)
    return need


def send_weight_update(named_parameters: Sequence[tuple[str, torch.Tensor]], group, bucket_bytes: int = 256 << 20,
                       src: int = 0, staging: _Staging | None = None, direct_bytes: int = 8 << 20) -> _Staging:
    """Sender side (finetune_loop.py:236-248): bucketed small tensors + direct large ones, then barrier."""
    params = [p.detach() for _, p in named_parameters]
    params = [p if p.is_contiguous() else p.contiguous() for p in params]
    infos = [ParameterInfo(name=n, shape=list(p.shape), dtype=str(torch.bfloat16)) for n, p in named_parameters]
    plan = plan_transfers(infos, bucket_bytes, direct_bytes)
    need = _staging_need(infos, plan)
    dev = params[0].device
    if staging is None or staging.bufs[0].numel() < need:
        staging = _Staging(dev, need)
    main = torch.cuda.current_stream(dev)
    ready = torch.cuda.Event()
    ready.record(main)  # parameters are final once the current stream reaches this point
    slot = 0
    for kind, item in plan:
        if kind == "direct" and params[item].dtype == torch.bfloat16:
            with torch.cuda.stream(staging.comm):  # zero-copy: broadcast straight out of the parameter storage
                staging.comm.wait_event(ready)
                dist.broadcast(params[item].view(-1), src=src, group=group)
            continue
        idxs = item if kind == "bucket" else [item]
        n = _bucket_elems(infos, idxs)
        buf = staging.bufs[slot][:n]
        main.wait_event(staging.free[slot])           # previous broadcast out of this slot finished
        ops.bucket_pack([params[i] for i in idxs], buf)  # K6 (fp32 masters are cast to bf16 while packing)
        packed = torch.cuda.Event()
        packed.record(main)
        with torch.cuda.stream(staging.comm):
            staging.comm.wait_event(packed)
            dist.broadcast(buf, src=src, group=group)
            staging.free[slot].record(staging.comm)
        slot ^= 1
    main.wait_stream(staging.comm)
    dist.barrier(group)
    return staging


def receive_weight_update(request: WeightUpdateRequest, group, load_weights: Callable[[list[tuple[str, torch.Tensor]]], object],
                          device, model_dtype: torch.dtype = torch.bfloat16, bucket_bytes: int = 256 << 20, src: int = 0,
                          staging: _Staging | None = None, direct_bytes: int = 8 << 20) -> _Staging:
    """Receiver side (vllm1.py:81-94): same checks, same plan, zero-copy views into the received buffers."""
    infos = request.parameters_info
    for info in infos:
        assert info.dtype == str(model_dtype), f"mismatch dtype: src {info.dtype}, dst {model_dtype}"
    plan = plan_transfers(infos, bucket_bytes, direct_bytes)
    need = _staging_need(infos, plan)
    if staging is None or staging.bufs[0].numel() < need:
        staging = _Staging(device, need)
    main = torch.cuda.current_stream(device)
    slot = 0
    for kind, item in plan:
        idxs = item if kind == "bucket" else [item]
        n = _bucket_elems(infos, idxs) if kind == "bucket" else _numel(infos[item].shape)
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
        slot ^= 1
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


# ---------------------------------------------------------------------------------------------------------
# NVSwitch multicast path: pack and broadcast are ONE kernel
# ---------------------------------------------------------------------------------------------------------
def plan_multicast(parameters_info: Sequence[ParameterInfo], bucket_bytes: int) -> list[list[int]]:
    """Transfers of the multicast path: in-order groups of parameters, each at most ``bucket_bytes`` of bf16
    (a larger tensor travels alone).  Both sides derive it from the request; identical to ``plan_buckets``."""
    return plan_buckets(parameters_info, bucket_bytes)


class MulticastChannel:
    """Two symmetric staging slots mapped through an NVSwitch multicast address (torch symmetric memory is the
    allocator / rendezvous; the data path is ``b200grpo_bucket_pack_multicast``).

    Per transfer k (slot k % 2) the sender runs ONE kernel that reads the parameters (bf16, or fp32 masters cast
    on the fly) and stores them to the multicast address -- the switch writes every receiver's copy of the slot --
    then all ranks meet in one barrier; receivers hand ``load_weights`` zero-copy views of their local slot.  A
    receiver reaches barrier k+1 only after its loads of transfer k were issued in stream order, so the sender's
    pack of transfer k+2 (same slot) cannot overtake them: one barrier per transfer, pack k+1 overlaps load k.

    Every rank of ``group`` must construct the channel collectively with the same ``slot_bytes``.
    """

    def __init__(self, group, device, slot_bytes: int):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.device = torch.device(device)
        self.slot_elems = (int(slot_bytes) // 2 + 7) // 8 * 8
        self.buf = symm.empty(2 * self.slot_elems, dtype=torch.bfloat16, device=self.device)
        self.handle = symm.rendezvous(self.buf, self.group)
        if not self.handle.has_multicast_support or not self.handle.multicast_ptr:
            raise RuntimeError("MulticastChannel: this group has no NVSwitch multicast mapping; use send/receive_weight_update")
        self.mc_ptr = int(self.handle.multicast_ptr)

    @staticmethod
    def slot_bytes_for(parameters_info: Sequence[ParameterInfo], bucket_bytes: int) -> int:
        plan = plan_multicast(parameters_info, bucket_bytes)
        return 2 * max(_bucket_elems(parameters_info, idxs) for idxs in plan)

    def send(self, named_parameters: Sequence[tuple[str, torch.Tensor]], bucket_bytes: int = 256 << 20) -> int:
        """Sender (finetune_loop.py:236-248).  Returns the number of transfers."""
        params = [p.detach() for _, p in named_parameters]
        params = [p if p.is_contiguous() else p.contiguous() for p in params]
        infos = [ParameterInfo(name=n, shape=list(p.shape), dtype=str(torch.bfloat16)) for n, p in named_parameters]
        plan = plan_multicast(infos, bucket_bytes)
        for k, idxs in enumerate(plan):
            n = _bucket_elems(infos, idxs)
            if n > self.slot_elems:
                raise ValueError(f"transfer {k} needs {2 * n} bytes, slots hold {2 * self.slot_elems}")
            ops.bucket_pack_multicast([params[i] for i in idxs], self.mc_ptr + 2 * (k % 2) * self.slot_elems, n)
            self.handle.barrier(channel=0)
        self.handle.barrier(channel=0)  # receivers finished loading the last transfers
        return len(plan)

    def receive(self, request: WeightUpdateRequest, load_weights: Callable[[list[tuple[str, torch.Tensor]]], object],
                model_dtype: torch.dtype = torch.bfloat16, bucket_bytes: int = 256 << 20) -> int:
        """Receiver (vllm1.py:81-94): same checks as the reference, views into the local copy of the slot."""
        infos = request.parameters_info
        for info in infos:
            assert info.dtype == str(model_dtype), f"mismatch dtype: src {info.dtype}, dst {model_dtype}"
        plan = plan_multicast(infos, bucket_bytes)
        for k, idxs in enumerate(plan):
            self.handle.barrier(channel=0)      # transfer k has landed in this rank's slot
            base = (k % 2) * self.slot_elems
            off = 0
            for i in idxs:
                numel = _numel(infos[i].shape)
                view = self.buf[base + off: base + off + numel].view(tuple(infos[i].shape))
                loaded = load_weights([(infos[i].name, view)])
                if loaded is not None and len(loaded) != 1:
                    raise ValueError(f"model {infos[i].name} not found in model state dict")
                off += (numel + 7) // 8 * 8
        self.handle.barrier(channel=0)
        return len(plan)
