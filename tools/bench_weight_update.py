#!/usr/bin/env python
"""BASELINE config 5: in-flight weight broadcast of Qwen2.5-7B bf16 parameters, trainer rank 0 -> actor GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/bench_weight_update.py [--buckets-mb 1,4,16,64,256,1024] [--scatter-mb 16,64,256] [--iters 3] [--fp32-masters]

Rank 0 plays the trainer (finetune_loop.py:236-248), ranks 1..N-1 the vLLM workers (vllm1.py:81-94) whose
`load_weights` is modelled as a copy of each received tensor into the resident model parameter.  Times the
reference's per-parameter loop (restated in weight_update.reference_send/receive) and the bucketed path for
each bucket size of SURVEY 8d's sweep, the NVSwitch multicast path (pack + broadcast in one kernel) and the scatter + relay
path (pack-scatter over P2P stores, multicast relay by the receivers); checks byte equality of every received parameter; prints one JSON line per variant with
GB/s = bf16 parameter bytes delivered per receiver / max-over-ranks device time.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def qwen25_7b_shapes(layers=28, hidden=3584, inter=18944, vocab=152064, kv=512):
    """The 339 parameter tensors of Qwen2.5-7B in state-dict order."""
    s = [("model.embed_tokens.weight", (vocab, hidden))]
    for i in range(layers):
        p = f"model.layers.{i}."
        s += [(p + "self_attn.q_proj.weight", (hidden, hidden)), (p + "self_attn.q_proj.bias", (hidden,)),
              (p + "self_attn.k_proj.weight", (kv, hidden)), (p + "self_attn.k_proj.bias", (kv,)),
              (p + "self_attn.v_proj.weight", (kv, hidden)), (p + "self_attn.v_proj.bias", (kv,)),
              (p + "self_attn.o_proj.weight", (hidden, hidden)),
              (p + "mlp.gate_proj.weight", (inter, hidden)), (p + "mlp.up_proj.weight", (inter, hidden)),
              (p + "mlp.down_proj.weight", (hidden, inter)),
              (p + "input_layernorm.weight", (hidden,)), (p + "post_attention_layernorm.weight", (hidden,))]
    s += [("model.norm.weight", (hidden,)), ("lm_head.weight", (vocab, hidden))]
    return s


def run_sweep(dev, rank: int, world: int, buckets_mb=(1, 4, 16, 64, 256, 1024), multicast_mb=(64, 256), scatter_mb=(16, 64, 256),
              iters: int = 3, fp32_masters: bool = False, layers: int = 28, stripe_kb: int = 64, reference: bool = True,
              scatter_p2p_mb=None) -> list[dict]:
    """Every transport over the default process group (rank 0 = trainer, the others = actor GPUs); returns one record per
    variant (identical on every rank).  GB/s = bf16 parameter bytes delivered per receiver / max-over-ranks device time."""
    from pipelinerl_swe_b200 import weight_update as wu

    shapes = qwen25_7b_shapes(layers=layers)
    n_bytes = sum(2 * int(torch.Size(s).numel()) for _, s in shapes)
    src_dtype = torch.float32 if (fp32_masters and rank == 0) else torch.bfloat16
    params = []
    for i, (name, shape) in enumerate(shapes):  # arange-patterned values: byte-equality is checkable on every rank
        n = int(torch.Size(shape).numel())
        t = ((torch.arange(n, device=dev, dtype=torch.int32) * 31 + i * 7) % 251).to(torch.float32).sub_(125.0).mul_(1.0 / 64)
        params.append((name, t.to(src_dtype if rank == 0 else torch.bfloat16).reshape(shape)))
    expect = None
    if rank != 0:
        expect = [p.clone() for _, p in params]
        for _, p in params:
            p.zero_()
    by_name = {n: p for n, p in params}

    def load_weights(items):
        for name, t in items:
            by_name[name].copy_(t)
        return [n for n, _ in items]

    request = wu.make_request(params, version=1)

    def timed(fn):
        ts = []
        for _ in range(iters + 1):
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t))
        return min(ts[1:])  # first iteration warms NCCL channels / allocator

    def check():
        ok = True
        if rank != 0:
            ok = all(torch.equal(p.view(torch.int16), e.view(torch.int16)) for (_, p), e in zip(params, expect))
            for _, p in params:
                p.zero_()
        t = torch.tensor([int(ok)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    out = []

    def emit(variant, ms, ok, extra=None):
        line = {"metric": "weight_broadcast_bf16_GBps", "variant": variant, "value": n_bytes / ms / 1e6, "unit": "GB/s",
                "ms": ms, "n_gpus": world, "receivers": world - 1, "payload_GB": n_bytes / 1e9, "tensors": len(shapes),
                "byte_exact": ok, "fp32_masters": bool(fp32_masters)}
        line.update(extra or {})
        out.append(line)

    # reference behaviour: one broadcast (+ alloc + load) per parameter
    if reference:
        if rank == 0:
            ms = timed(lambda: wu.reference_send(params, None))
        else:
            ms = timed(lambda: wu.reference_receive(request, None, load_weights, dev))
        emit("reference_per_parameter", ms, check(), {"n_broadcasts": len(shapes)})

    for mb in buckets_mb:
        staging = [None]

        def run():
            if rank == 0:
                staging[0] = wu.send_weight_update(params, None, bucket_bytes=mb << 20, staging=staging[0])
            else:
                staging[0] = wu.receive_weight_update(request, None, load_weights, dev, bucket_bytes=mb << 20, staging=staging[0])

        ms = timed(run)
        emit(f"bucketed_{mb}MB", ms, check(), {"bucket_MB": mb, "n_broadcasts": len(wu.plan_transfers(request.parameters_info, mb << 20))})
        staging[0] = None
        torch.cuda.empty_cache()
    # NVSwitch multicast: pack + broadcast in one kernel; scatter + relay: pack-scatter over P2P, then multicast relay
    for kind, sizes in (("multicast", multicast_mb), ("scatter", scatter_mb),
                        ("scatter_p2p", scatter_mb if scatter_p2p_mb is None else scatter_p2p_mb)):
        for mb in sizes:
            try:
                if kind == "multicast":
                    ch = wu.MulticastChannel(None, dev, wu.MulticastChannel.slot_bytes_for(request.parameters_info, mb << 20))
                else:
                    ch = wu.ScatterChannel(None, dev, wu.ScatterChannel.slot_bytes_for(request.parameters_info, mb << 20),
                                           stripe_bytes=stripe_kb << 10, relay="p2p" if kind == "scatter_p2p" else "multicast")
            except Exception as e:  # noqa: BLE001
                out.append({"variant": f"{kind}_{mb}MB", "unavailable": repr(e)[:200]})
                break

            def run_ch():
                if rank == 0:
                    ch.send(params, bucket_bytes=mb << 20)
                else:
                    ch.receive(request, load_weights, bucket_bytes=mb << 20)

            ms = timed(run_ch)
            emit(f"{kind}_{mb}MB", ms, check(), {"bucket_MB": mb, "n_transfers": len(wu.plan_multicast(request.parameters_info, mb << 20)),
                                               "slot_MB": ch.slot_elems * 2 / 2 ** 20, **({"stripe_KB": stripe_kb} if kind.startswith("scatter") else {})})
            del ch
            torch.cuda.empty_cache()
    del params, expect, by_name
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--buckets-mb", default="1,4,16,64,256,1024")
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--fp32-masters", action="store_true", help="sender holds fp32 parameters (cast fused into the pack)")
    ap.add_argument("--layers", type=int, default=28)
    ap.add_argument("--multicast-mb", default="64,256", help="bucket sizes for the NVSwitch multicast path ('' to skip)")
    ap.add_argument("--scatter-mb", default="16,64,256", help="bucket sizes for the scatter + relay path ('' to skip)")
    ap.add_argument("--stripe-kb", type=int, default=64)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ints = lambda s: tuple(int(x) for x in s.split(",") if x)  # noqa: E731
    for line in run_sweep(dev, rank, world, ints(args.buckets_mb), ints(args.multicast_mb), ints(args.scatter_mb), args.iters,
                          args.fp32_masters, args.layers, args.stripe_kb):
        if rank == 0:
            print(json.dumps(line), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
