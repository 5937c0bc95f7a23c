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
                       src: int = 0, staging: _Staging | None = None, direct_bytes: int = 8 << 20, gather=None,
                       is_main: bool = True, final_barrier: bool = True, shapes: Sequence[Sequence[int]] | None = None) -> _Staging | None:
    """Sender side (finetune_loop.py:236-248): bucketed small tensors + direct large ones, then barrier.

    ``gather`` (ZeRO-3, finetune_loop.py:202-205): a context-manager factory; the parameters of ONE transfer of the plan are
    whole only inside ``with gather(those_parameters):`` -- a bucket's worth of small tensors or one large tensor at a time,
    never the whole model -- and every trainer rank enters the contexts in plan order while only ``is_main`` packs and
    broadcasts.  ``shapes``: the full shapes when the parameters are sharded outside the contexts (``ds_shape``)."""
    named_parameters = list(named_parameters)
    params = [p.detach() for _, p in named_parameters]
    shapes = [list(sh) for sh in shapes] if shapes is not None else [list(p.shape) for p in params]
    infos = [ParameterInfo(name=n, shape=sh, dtype=str(torch.bfloat16)) for (n, _), sh in zip(named_parameters, shapes)]
    plan = plan_transfers(infos, bucket_bytes, direct_bytes)
    need = _staging_need(infos, plan)
    dev = params[0].device
    if is_main and (staging is None or staging.bufs[0].numel() < need):
        staging = _Staging(dev, need)
    main = torch.cuda.current_stream(dev)
    ready = torch.cuda.Event()
    ready.record(main)  # parameters are final once the current stream reaches this point
    slot = 0

    def whole(i):  # the parameter as it is NOW (inside a gather context: its full value)
        t = named_parameters[i][1].detach()
        return t if t.is_contiguous() else t.contiguous()

    for kind, item in plan:
        idxs = item if kind == "bucket" else [item]
        ctx = gather([named_parameters[i][1] for i in idxs]) if gather is not None else None
        if ctx is not None:
            ctx.__enter__()
        try:
            if not is_main:
                continue
            if gather is not None:
                ready.record(main)
            if kind == "direct" and whole(item).dtype == torch.bfloat16:
                with torch.cuda.stream(staging.comm):  # zero-copy: broadcast straight out of the parameter storage
                    staging.comm.wait_event(ready)
                    dist.broadcast(whole(item).view(-1), src=src, group=group)
                continue
            n = _bucket_elems(infos, idxs)
            buf = staging.bufs[slot][:n]
            main.wait_event(staging.free[slot])           # previous broadcast out of this slot finished
            ops.bucket_pack([whole(i) for i in idxs], buf)  # K6 (fp32 masters are cast to bf16 while packing)
            packed = torch.cuda.Event()
            packed.record(main)
            # what crosses the wire: a bucket with its 16-byte padding, a direct tensor (fp32 master cast through the staging
            # slot) WITHOUT it -- exactly the element counts receive_weight_update posts
            wire = buf if kind == "bucket" else buf[:_numel(infos[item].shape)]
            with torch.cuda.stream(staging.comm):
                staging.comm.wait_event(packed)
                dist.broadcast(wire, src=src, group=group)
                staging.free[slot].record(staging.comm)
            slot ^= 1
        finally:
            if ctx is not None:
                if is_main:
                    main.wait_stream(staging.comm)        # the gathered storage is released in main-stream order: after its broadcast
                ctx.__exit__(None, None, None)
    if is_main:
        main.wait_stream(staging.comm)
        if final_barrier:
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


# ---------------------------------------------------------------------------------------------------------
# Scatter + relay: 1 -> N at the line rate of one copy
# ---------------------------------------------------------------------------------------------------------
class ScatterChannel:
    """The broadcast as two kernels of this library over a symmetric three-slot buffer (torch symmetric memory is only the
    allocator / rendezvous / barrier): the sender's pack kernel stores stripe j of the bucket straight into receiver
    (j mod N)'s slot over P2P NVLink (``b200grpo_bucket_pack_scatter``: parameters read once, no staging copy, egress =
    the payload once), then each receiver re-emits its stripes with ``multimem.st`` (``b200grpo_bucket_relay_multicast``)
    and the switch completes every other copy (receiver ingress = the payload once, egress = payload / N).  NCCL's
    broadcast of the same bucket pushes the payload down a chain of GPUs; here no link carries it twice.

    Stream schedule (three slots, two barrier channels A / B; transfer k uses slot k % 3):
        sender          : scatter(0) A(0) scatter(1) B(0) A(1) scatter(2) B(1) A(2) ...    (scatter(k+1) overlaps relay(k))
        receiver, main  : [wait load(k-2)] A(k) relay(k) B(k)                               for k = 0, 1, ...
        receiver, side  : load(k) after B(k)                                                (load(k) overlaps relay(k+1))
    scatter(k+3) reuses slot k % 3 and is issued after A(k+2), which a receiver enters only after load(k) has finished.
    With two slots and the loads on the main stream the receivers' relay -> load chain was serial (0.63 ms per 256 MB
    transfer against 0.46 ms of pure transfer time).

    Every rank of ``group`` constructs the channel collectively; rank ``src`` is the sender, the others are receivers."""

    def __init__(self, group, device, slot_bytes: int, src: int = 0, stripe_bytes: int = 64 << 10, relay_ctas: int = 0,
                 relay: str = "multicast"):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world, self.src = dist.get_rank(self.group), dist.get_world_size(self.group), src
        assert self.world >= 2, "ScatterChannel needs a sender and at least one receiver"
        self.device = torch.device(device)
        self.slot_elems = (int(slot_bytes) // 2 + 7) // 8 * 8
        self.stripe_elems = max(8, int(stripe_bytes) // 2 // 8 * 8)
        self.relay_ctas, self.relay = relay_ctas, relay          # relay: "multicast" (multimem.st) or "p2p" (plain stores to each peer)
        self.buf = symm.empty(self.SLOTS * self.slot_elems, dtype=torch.bfloat16, device=self.device)
        self.handle = symm.rendezvous(self.buf, self.group)
        self.side = torch.cuda.Stream(self.device)
        if relay == "multicast" and (not self.handle.has_multicast_support or not self.handle.multicast_ptr):
            raise RuntimeError("ScatterChannel: this group has no NVSwitch multicast mapping; use relay='p2p' or send/receive_weight_update")
        self.mc_ptr = int(self.handle.multicast_ptr) if relay == "multicast" else 0
        self.receivers = [r for r in range(self.world) if r != src]            # peer index q -> rank
        self.peer_ptrs = [int(self.handle.buffer_ptrs[r]) for r in self.receivers]
        self.peer_index = self.receivers.index(self.rank) if self.rank != src else -1

    slot_bytes_for = staticmethod(MulticastChannel.slot_bytes_for)
    SLOTS = 3

    def send(self, named_parameters: Sequence[tuple[str, torch.Tensor]], bucket_bytes: int = 64 << 20) -> int:
        """Sender (finetune_loop.py:236-248).  Returns the number of transfers."""
        assert self.rank == self.src
        params = [p.detach() for _, p in named_parameters]
        params = [p if p.is_contiguous() else p.contiguous() for p in params]
        infos = [ParameterInfo(name=n, shape=list(p.shape), dtype=str(torch.bfloat16)) for n, p in named_parameters]
        plan = plan_multicast(infos, bucket_bytes)
        for k, idxs in enumerate(plan):
            n = _bucket_elems(infos, idxs)
            if n > self.slot_elems:
                raise ValueError(f"transfer {k} needs {2 * n} bytes, slots hold {2 * self.slot_elems}")
            base = 2 * (k % self.SLOTS) * self.slot_elems               # byte offset of the slot
            ops.bucket_pack_scatter([params[i] for i in idxs], [p + base for p in self.peer_ptrs], self.stripe_elems, n)
            if k >= 1:
                self.handle.barrier(channel=1)                          # B(k-1): relays of transfer k-1 have landed everywhere
            self.handle.barrier(channel=0)                              # A(k): stripes of transfer k are in their owners' slots
        if plan:
            self.handle.barrier(channel=1)                              # B(K-1)
            self.handle.barrier(channel=0)                              # every receiver has finished its loads: the slots are free for the next update
        return len(plan)

    def receive(self, request: WeightUpdateRequest, load_weights: Callable[[list[tuple[str, torch.Tensor]]], object],
                model_dtype: torch.dtype = torch.bfloat16, bucket_bytes: int = 64 << 20) -> int:
        """Receiver (vllm1.py:81-94): same checks as the reference, views into the local copy of the slot."""
        assert self.rank != self.src
        infos = request.parameters_info
        for info in infos:
            assert info.dtype == str(model_dtype), f"mismatch dtype: src {info.dtype}, dst {model_dtype}"
        plan = plan_multicast(infos, bucket_bytes)
        local_ptr = int(self.handle.buffer_ptrs[self.rank])
        main = torch.cuda.current_stream(self.device)
        loaded: list[torch.cuda.Event] = []
        for k, idxs in enumerate(plan):
            n = _bucket_elems(infos, idxs)
            base = (k % self.SLOTS) * self.slot_elems
            if k >= 2:
                main.wait_event(loaded[k - 2])                          # the sender may refill that slot once everybody passes A(k)
            self.handle.barrier(channel=0)                              # A(k): my stripes of transfer k have arrived
            if self.relay == "multicast":
                ops.bucket_relay_multicast(local_ptr + 2 * base, self.mc_ptr + 2 * base, n, self.peer_index, len(self.receivers),
                                           self.stripe_elems, self.relay_ctas)
            else:
                ops.bucket_relay_p2p(local_ptr + 2 * base, [p + 2 * base for p in self.peer_ptrs], n, self.peer_index,
                                     self.stripe_elems, self.relay_ctas)
            self.handle.barrier(channel=1)                              # B(k): everybody's stripes are in my slot
            landed = torch.cuda.Event()
            landed.record(main)
            with torch.cuda.stream(self.side):                          # load(k) beside relay(k+1)
                self.side.wait_event(landed)
                off = 0
                for i in idxs:
                    numel = _numel(infos[i].shape)
                    view = self.buf[base + off: base + off + numel].view(tuple(infos[i].shape))
                    loaded_names = load_weights([(infos[i].name, view)])
                    if loaded_names is not None and len(loaded_names) != 1:
                        raise ValueError(f"model {infos[i].name} not found in model state dict")
                    off += (numel + 7) // 8 * 8
                done = torch.cuda.Event()
                done.record(self.side)
            loaded.append(done)
        main.wait_stream(self.side)
        if plan:
            self.handle.barrier(channel=0)                              # tells the sender the slots are free again
        return len(plan)


# ---------------------------------------------------------------------------------------------------------
# The second NCCL world and the two call shapes of the reference
# ---------------------------------------------------------------------------------------------------------
def init_actor_update_group(init_method: str, rank: int, world_size: int, group_name: str = "actor", backend: str = "nccl",
                            timeout=None, device: torch.device | None = None):
    """A process group that lives BESIDE the default one: trainer rank 0 (``rank`` 0) + every actor GPU
    (``1 + actor_idx * actor_ngpus + local rank``), rendezvousing at ``init_method`` (tcp://host:port) -- the role of
    /root/reference/pipelinerl/torch_utils.py:16-65 ``init_extra_process_group`` as called from
    finetune_loop.py (trainer) and vllm1.py:52-79 (workers).  ``torch.distributed.init_process_group`` cannot be used: each
    side already has (or will have) its own default group.  The group is built on a PrefixStore named after the group, so
    its keys cannot collide with another world's on the same TCP store."""
    from datetime import timedelta
    from torch.distributed import distributed_c10d as c10d
    if not init_method:
        raise ValueError("init_actor_update_group needs an explicit init_method (tcp://...): the default env:// rendezvous "
                         "belongs to the process's own world")
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside a world of {world_size}")
    timeout = timeout if timeout is not None else c10d.default_pg_timeout
    if not isinstance(timeout, timedelta):
        timeout = timedelta(seconds=float(timeout))
    if init_method.startswith("tcp://"):
        # Host the store explicitly on group rank 0.  torch's own tcp:// rendezvous handler makes EVERY rank a client of the
        # launcher's agent store when the process runs under torchrun / accelerate launch (TORCHELASTIC_USE_AGENT_STORE), and
        # nobody then listens on this second world's port.
        from urllib.parse import urlparse
        url = urlparse(init_method)
        store = dist.TCPStore(url.hostname, int(url.port), world_size, rank == 0, timeout, wait_for_workers=True, multi_tenant=True)
    else:
        store, rank, world_size = next(c10d.rendezvous(init_method, rank, world_size, timeout=timeout))
    store.set_timeout(timeout)
    kw = {}
    if (device is not None and torch.device(device).type == "cuda" and torch.device(device).index is not None
            and "device_id" in c10d._new_process_group_helper.__code__.co_varnames):
        kw["device_id"] = torch.device(device)   # binds the NCCL communicator to this GPU (eager init, no device guessing in barrier)
    pg, _ = c10d._new_process_group_helper(world_size, rank, [], c10d.Backend(backend), c10d.PrefixStore(group_name, store),
                                           group_name=group_name, timeout=timeout, **kw)
    c10d._world.pg_group_ranks[pg] = {r: r for r in range(world_size)}   # global rank == group rank in this world
    return pg


class ActorUpdateWorker:
    """Receiver side with the reference's method shapes (vllm1.py ``WorkerExtension``, :47-94): mix into an object that has
    ``rank``, ``device``, ``model_config.dtype`` and ``model_runner.model.load_weights``.

    ``transport``: "nccl" = bucketed broadcasts (``receive_weight_update`` above), "scatter" = ``ScatterChannel``,
    "reference" = the reference's own per-parameter loop."""

    weight_update_transport: str = "nccl"
    weight_update_bucket_bytes: int = 256 << 20

    def init_actor_update_group(self, actor_idx: int, actor_ngpus: int, weight_update_group_init_method: str,
                                weight_update_group_world_size: int):
        self.pg_rank = 1 + actor_idx * actor_ngpus + self.rank                       # vllm1.py:61
        self.process_group = init_actor_update_group(weight_update_group_init_method, self.pg_rank, weight_update_group_world_size,
                                                     device=getattr(self, "device", None))
        self._wu_state = None
        return self.process_group

    def receive_weight_update(self, request: WeightUpdateRequest):
        if torch.device(self.device).type == "cuda":
            torch.cuda.synchronize(self.device)                                       # vllm1.py:82
        load = lambda weights: self.model_runner.model.load_weights(weights=weights)  # noqa: E731
        dtype = self.model_config.dtype
        if self.weight_update_transport == "reference":
            for info in request.parameters_info:
                assert info.dtype == str(dtype), f"mismatch dtype: src {info.dtype}, dst {dtype}"
            reference_receive(request, self.process_group, load, self.device)
        elif self.weight_update_transport == "scatter":
            if self._wu_state is None:
                self._wu_state = ScatterChannel(self.process_group, self.device,
                                                ScatterChannel.slot_bytes_for(request.parameters_info, self.weight_update_bucket_bytes))
            self._wu_state.receive(request, load, dtype, self.weight_update_bucket_bytes)
        else:
            self._wu_state = receive_weight_update(request, self.process_group, load, self.device, dtype,
                                                   self.weight_update_bucket_bytes, staging=self._wu_state)


class WeightUpdateSender:
    """Trainer side with the reference's call shape, ``send_weight_update(version)`` (finetune_loop.py:174-256).

    ``model``: the (unwrapped) module whose parameters are sent; ``publish(request)``: what the reference does with the
    ``WeightUpdateRequest`` (HTTP POST to every actor, :160-172) -- returns an iterable of futures or None;
    ``gather``: for ZeRO-3 style sharded parameters a context-manager factory ``gather(params)`` inside which the given
    parameters hold their full values (``deepspeed.zero.GatheredParameters``, :202-205); the parameter's full shape is
    read from ``ds_shape`` when present.  Only ``is_main`` ranks broadcast; every trainer rank must call
    ``send_weight_update`` so that all of them enter the gathers.  (deepspeed is not in this image: the branch is tested
    with a stand-in context manager over sharded tensors, tests/test_multirank_cpu.py.)"""

    def __init__(self, model: torch.nn.Module, actor_update_group, publish: Callable[[WeightUpdateRequest], object] | None = None,
                 transport: str = "nccl", bucket_bytes: int = 256 << 20, gather=None, is_main: bool = True,
                 gather_group_bytes: int = 256 << 20):
        self.model, self.group, self.publish = model, actor_update_group, publish
        self.transport, self.bucket_bytes, self.gather, self.is_main = transport, bucket_bytes, gather, is_main
        self.gather_group_bytes = gather_group_bytes
        self._state = None

    def _named_parameters(self) -> list[tuple[str, torch.Tensor]]:
        inner = getattr(self.model, "pretrained_model", self.model)     # value-head wrapper: only the LM is sent (:186-190)
        return list(inner.named_parameters())

    def _send(self, named: list[tuple[str, torch.Tensor]]) -> None:
        if self.transport == "reference":
            for _, p in named:
                dist.broadcast(p.data.bfloat16(), src=0, group=self.group)
        elif self.transport == "scatter":
            if self._state is None:
                infos = [ParameterInfo(name=n, shape=list(p.shape), dtype=str(torch.bfloat16)) for n, p in named]
                self._state = ScatterChannel(self.group, named[0][1].device, ScatterChannel.slot_bytes_for(infos, self.bucket_bytes))
            self._state.send(named, self.bucket_bytes)
        else:
            self._state = send_weight_update(named, self.group, self.bucket_bytes, staging=self._state, final_barrier=False)

    def send_weight_update(self, version: int) -> WeightUpdateRequest:
        named = self._named_parameters()
        shape = lambda p: list(getattr(p, "ds_shape", p.shape))  # noqa: E731   (:193-196: ZeRO-3 parameters know their full shape)
        request = WeightUpdateRequest(version=version, parameters_info=[
            ParameterInfo(name=n, shape=shape(p), dtype=str(torch.bfloat16)) for n, p in named])
        futures = self.publish(request) if (self.is_main and self.publish is not None) else None
        if self.gather is None:
            if self.is_main:
                self._send(named)
        elif self.transport == "reference":   # the reference's own loop: ONE parameter gathered per collective (:202-205)
            for n, p in named:
                with self.gather([p]):
                    if self.is_main:
                        dist.broadcast(p.data.bfloat16(), src=0, group=self.group)
        elif self.transport == "nccl":        # one transfer of the plan gathered at a time, packed and broadcast while whole
            self._state = send_weight_update(named, self.group, self.bucket_bytes, staging=self._state, gather=self.gather,
                                             is_main=self.is_main, final_barrier=False, shapes=[shape(p) for _, p in named])
        else:
            raise ValueError("gathered (ZeRO-3) parameters stream through transport 'nccl' or 'reference'")
        if self.is_main:
            if self.transport != "scatter":
                dist.barrier(self.group)
            for f in (futures or []):
                f.result()
        return request
