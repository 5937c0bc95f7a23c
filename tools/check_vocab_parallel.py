#!/usr/bin/env python
"""Vocab-parallel logprob/loss over NCCL == unsharded, on real GPUs (run under torchrun, N = 2/4/8).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/check_vocab_parallel.py [--exchange=symm]

Every rank holds the same hidden/targets and its contiguous shard of lm_head rows; forward exchanges the
16 B/token partials with one all-gather, backward all-reduces dH.  Compared on every rank against the
unsharded kernels run locally on the full weight: logprob/entropy bit-identical across ranks and within
1e-4 of unsharded, dW shard within 1e-2 (bf16) of the matching rows, dH within 2e-2 (the north star's gradient
tolerance: NCCL's ring sums the 8 bf16 partials in bf16, 1.4e-2 at 8 ranks; the multicast kernel accumulates in f32, 4.5e-3).
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
    ok = identical and e_out < 1e-4 and e_dw < 1e-2 and e_dh < 2e-2   # dH: the north star's 2e-2 (NCCL sums 8 bf16 partials in bf16)

    # ---- the drop-in itself: rl_step(model, batch, step, max_step, config, vocab_group=...) on a model whose lm_head holds
    #      only this rank's vocab rows == rl_step on the unsharded model (loss, all stats, dH, this rank's rows of dW)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import StubLM, synthetic_packed_batch
    from pipelinerl_swe_b200 import RLConfig, rl_step
    Hs, Vs, lens = 256, 32000 + 8, [300, 140, 260]
    batch = synthetic_packed_batch(lens, Vs, seed=5)
    dbatch = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
    cfg = RLConfig(policy_loss="ppo", batch_size=3, kl_coef=0.1, final_kl_coef=0.1, entropy_bonus=0.01, final_entropy_bonus=0.01,
                   temperature=0.9)
    full = StubLM(Hs, Vs, sum(lens), seed=8).to(dev)
    loss_f, stats_f = rl_step(full, dbatch, 1, 10, cfg)
    loss_f.backward()
    shard = StubLM(Hs, Vs, sum(lens), seed=8).to(dev)
    w_sh, off = shard_weight(shard.lm_head.weight.detach(), world, rank)
    shard.lm_head = torch.nn.Linear(Hs, w_sh.shape[0], bias=False, dtype=torch.bfloat16, device=dev)
    with torch.no_grad():
        shard.lm_head.weight.copy_(w_sh)
    group = SymmetricExchange(dist.group.WORLD, dev, sum(lens), Hs) if symmetric else dist.group.WORLD
    loss_s, stats_s = rl_step(shard, dbatch, 1, 10, cfg, vocab_group=group)
    loss_s.backward()
    assert shard.lm_head.v_offset == off, (shard.lm_head.v_offset, off)
    e_loss = abs(loss_s.item() - loss_f.item())
    e_stats = max(abs(stats_s[k] - stats_f[k]) / (abs(stats_f[k]) + 0.1) for k in stats_f)   # 1e-3 relative with a 1e-4 absolute floor
    e_rl_dh = rel(shard.hidden.grad, full.hidden.grad)
    e_rl_dw = rel(shard.lm_head.weight.grad, full.lm_head.weight.grad[off:off + w_sh.shape[0]])
    rl_ok = set(stats_s) == set(stats_f) and e_loss < 1e-5 + 1e-3 * abs(loss_f.item()) and e_stats < 1e-3 and e_rl_dh < 2e-2 and e_rl_dw < 1e-2
    ok = ok and rl_ok
    t = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"VOCAB_PARALLEL transport={'nvswitch-multicast' if symmetric else 'nccl'} world={world} {'OK' if t.item() else 'FAIL'} identical_across_ranks={identical} "
              f"max|d(logprob,entropy)|={e_out:.2e} rel dW={e_dw:.2e} rel dH={e_dh:.2e} | rl_step(vocab_group=...): "
              f"|d loss|={e_loss:.1e} stats rel={e_stats:.1e} rel dH={e_rl_dh:.2e} rel dW={e_rl_dw:.2e}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if t.item() else 1)


if __name__ == "__main__":
    main()
