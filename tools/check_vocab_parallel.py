#!/usr/bin/env python
"""Vocab-parallel logprob/loss over NCCL == unsharded, on real GPUs (run under torchrun, N = 2/4/8).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/check_vocab_parallel.py [--exchange=symm]

Every rank holds the same hidden/targets and its contiguous shard of lm_head rows; forward exchanges the
16 B/token partials with one all-gather, backward all-reduces dH.  Compared on every rank against the
unsharded kernels run locally on the full weight: logprob/entropy bit-identical across ranks and within
1e-4 of unsharded, dW shard within 1e-2 (bf16) of the matching rows, dH within 1e-2.
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from pipelinerl_swe_b200.lm_head import LMHeadLogprobFunction
    from pipelinerl_swe_b200.vocab_parallel import SymmetricExchange, shard_weight
    symmetric = "--exchange=symm" in sys.argv or os.environ.get("B200GRPO_EXCHANGE") == "symm"

    M, H, V, temp = 1000, 512, 152064, 0.9
    g = torch.Generator(device=dev).manual_seed(7)  # same seed on every rank -> same tensors
    hidden = torch.randn(M, H, device=dev, generator=g).bfloat16()
    weight = (torch.randn(V, H, device=dev, generator=g) * 0.04).bfloat16()
    targets = torch.randint(0, V, (M,), device=dev, generator=g)
    c = torch.randn(M, 2, device=dev, generator=g) * torch.tensor([0.1, 0.01], device=dev)

    h0, w0 = hidden.clone().requires_grad_(True), weight.clone().requires_grad_(True)
    out0 = LMHeadLogprobFunction.apply(h0, w0, targets, 1.0 / temp, None, 0)
    (out0 * c).sum().backward()

    w_shard, v_off = shard_weight(weight, world, rank)
    h1, w1 = hidden.clone().requires_grad_(True), w_shard.clone().requires_grad_(True)
    # transport: NCCL (the process group) or NVSwitch multicast kernels of this library (SymmetricExchange)
    transport = SymmetricExchange(dist.group.WORLD, dev, M, H) if symmetric else dist.group.WORLD
    out1 = LMHeadLogprobFunction.apply(h1, w1, targets, 1.0 / temp, transport, v_off)
    (out1 * c).sum().backward()
    torch.cuda.synchronize()
    if symmetric:   # a second micro-batch through the same buffers (other stats slot, reused dH buffer) must give the same result
        h2, w2 = hidden.clone().requires_grad_(True), w_shard.clone().requires_grad_(True)
        out2 = LMHeadLogprobFunction.apply(h2, w2, targets, 1.0 / temp, transport, v_off)
        (out2 * c).sum().backward()
        torch.cuda.synchronize()
        assert torch.equal(out2, out1) and torch.equal(h2.grad, h1.grad) and torch.equal(w2.grad, w1.grad), "second pass differs"

    def rel(a, b):
        return float((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-20))

    gathered = [torch.empty_like(out1) for _ in range(world)]
    dist.all_gather(gathered, out1.detach().contiguous())
    identical = all(torch.equal(gathered[0], x) for x in gathered)
    e_out = float((out1 - out0).abs().max())
    e_dw = rel(w1.grad, w0.grad[v_off:v_off + w_shard.shape[0]])
    e_dh = rel(h1.grad, h0.grad)
    ok = identical and e_out < 1e-4 and e_dw < 1e-2 and e_dh < 1e-2
    t = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"VOCAB_PARALLEL transport={'nvswitch-multicast' if symmetric else 'nccl'} world={world} {'OK' if t.item() else 'FAIL'} identical_across_ranks={identical} "
              f"max|d(logprob,entropy)|={e_out:.2e} rel dW={e_dw:.2e} rel dH={e_dh:.2e}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if t.item() else 1)


if __name__ == "__main__":
    main()
