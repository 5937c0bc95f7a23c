#!/usr/bin/env python
"""Trainer -> actor weight update through the reference's call shapes, every transport, byte-exact (run under torchrun, N >= 2).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 \
        tools/check_weight_update.py

Rank 0 is the trainer (``WeightUpdateSender.send_weight_update(version)``, finetune_loop.py:174-256), ranks 1..N-1 are vLLM
workers (``ActorUpdateWorker.init_actor_update_group`` / ``receive_weight_update(request)``, vllm1.py:52-94).  The update
group is a SECOND NCCL world built with ``init_actor_update_group`` over its own TCP rendezvous beside the default group
(torch_utils.py:16-65).  Parameters include fp32 masters, a large fp32 tensor with an odd element count (the 'direct' path
whose sender / receiver element counts must agree), odd-sized bf16 tensors and a ZeRO-3 stand-in (sharded storage, whole
only inside ``gather``).  Every received tensor must equal ``param.data.bfloat16()`` bit for bit.
"""
import contextlib
import os
import sys
import types

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import faulthandler
    faulthandler.dump_traceback_later(int(os.environ.get("WU_HANG_DUMP_S", "120")), exit=True)   # a hung collective names its line
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)                      # each side's own default world (bookkeeping only)
    from pipelinerl_swe_b200 import weight_update as wu

    port = int(os.environ.get("MASTER_PORT", "29500")) + 17
    init_method = f"tcp://127.0.0.1:{port}"

    # ---- the model on the trainer, the resident copy on the workers -------------------------------------------------
    g = torch.Generator(device=dev).manual_seed(5)                      # same values on every rank: receivers know the truth
    spec = [("embed.weight", (4099, 1024), torch.float32),              # large fp32 master, odd numel -> 'direct' through staging
            ("layers.0.q.weight", (1024, 1024), torch.bfloat16), ("layers.0.q.bias", (1021,), torch.bfloat16),
            ("layers.0.k.weight", (128, 1024), torch.float32), ("layers.0.norm.weight", (1024,), torch.float32),
            ("layers.0.mlp.weight", (5000, 1023), torch.bfloat16),      # large bf16, odd numel -> zero-copy 'direct'
            ("layers.1.q.weight", (1024, 1024), torch.bfloat16), ("layers.1.norm.weight", (7,), torch.bfloat16),
            ("lm_head.weight", (3001, 1024), torch.bfloat16)]
    full = {n: (torch.randn(*sh, device=dev, generator=g) * 0.5).to(dt) for n, sh, dt in spec}
    truth = {n: t.bfloat16() for n, t in full.items()}

    class Model(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.p = torch.nn.ParameterDict({n.replace(".", "__"): torch.nn.Parameter(t.clone(), requires_grad=False) for n, t in full.items()})

        def named_parameters(self, *a, **k):
            return [(n.replace("__", "."), p) for n, p in self.p.items()]

    results = []

    def run(transport, bucket_bytes, zero3=False):
        if rank == 0:
            model = Model()
            gather = None
            if zero3:   # ZeRO-3 stand-in: outside gather() a parameter holds only its first half; ds_shape is the full shape
                for n, p in model.named_parameters():
                    p.ds_shape = p.shape
                    p._full = p.data
                    p.data = p.data.reshape(-1)[: p.numel() // 2].clone()

                @contextlib.contextmanager
                def gather(params):
                    shards = [p.data for p in params]
                    for p in params:
                        p.data = p._full
                    try:
                        yield
                    finally:
                        for p, sh in zip(params, shards):
                            p.data = sh
            sender = wu.WeightUpdateSender(model, group, publish=None, transport=transport, bucket_bytes=bucket_bytes, gather=gather)
            sender.send_weight_update(version=7)

    # the update group: rank 0 = trainer, 1.. = actor GPUs (actor_idx 0, actor_ngpus world-1)
    group = None
    if rank == 0:
        group = wu.init_actor_update_group(init_method, 0, world, device=dev)   # blocks until every worker has joined the rendezvous

    class Worker(wu.ActorUpdateWorker):
        def __init__(self):
            self.rank, self.device = rank - 1, dev
            self.model_config = types.SimpleNamespace(dtype=torch.bfloat16)
            self.received = {}
            outer = self

            class M:
                def load_weights(self, weights):
                    for n, t in weights:
                        outer.received[n] = t.clone()
                    return [n for n, _ in weights]
            self.model_runner = types.SimpleNamespace(model=M())

    if rank != 0:
        worker = Worker()
        worker.init_actor_update_group(0, world - 1, init_method, world)
        assert worker.pg_rank == rank

    ok_all = True
    for transport, bucket_bytes, zero3 in (("reference", 0, False), ("nccl", 1 << 20, False), ("nccl", 64 << 20, False),
                                           ("scatter", 4 << 20, False), ("scatter", 64 << 20, False), ("nccl", 1 << 20, True),
                                           ("reference", 0, True)):
        # request first (default world stands in for the HTTP channel), then the data over the update group
        if rank == 0:
            model = Model()
            req = wu.make_request(model.named_parameters(), version=7)
            box = [req]
        else:
            box = [None]
        dist.broadcast_object_list(box, src=0)
        print(f"[rank {rank}] {transport} zero3={zero3} bucket={bucket_bytes}", file=sys.stderr, flush=True)
        ok = True
        try:
            if rank == 0:
                run(transport, bucket_bytes, zero3)
            else:
                worker.received.clear()
                worker.weight_update_transport, worker.weight_update_bucket_bytes = transport, max(bucket_bytes, 1 << 20)
                if transport == "scatter" and getattr(worker, "_scatter_bytes", None) != bucket_bytes:
                    worker._wu_state, worker._scatter_bytes = None, bucket_bytes
                if transport == "nccl":
                    worker._wu_state = None
                worker.receive_weight_update(box[0])
                torch.cuda.synchronize()
                ok = set(worker.received) == set(truth) and all(
                    torch.equal(worker.received[n].view(torch.int16), truth[n].view(torch.int16)) for n in truth)
        except Exception as e:  # noqa: BLE001
            import traceback
            traceback.print_exc()
            ok = False
        t = torch.tensor([int(ok)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        ok_all = ok_all and bool(t.item())
        results.append(f"{transport}{'+zero3' if zero3 else ''}@{bucket_bytes >> 20}MB={'ok' if t.item() else 'FAIL'}")
    if rank == 0:
        print(f"WEIGHT_UPDATE world={world} receivers={world - 1} {'OK' if ok_all else 'FAIL'} second-world=init_actor_update_group(tcp) "
              + " ".join(results), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()
