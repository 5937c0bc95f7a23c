#!/usr/bin/env python
"""bench.py -- GRPO logprob+loss fwd/bwd tokens/sec on B200 (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c3] [--impl reference]

One "step" = one packed micro-batch (T tokens) through the hot path: K1 lm_head GEMM + online
softmax/entropy/gather -> finalize -> K3 token loss (+34 metrics) -> K2 backward (dH, dW accumulated
into one f32 .grad).  Under torchrun (N>1) every rank runs an independent replica of the workload on its
own GPU (micro-batches are independent units; no data-path collective) and rank 0 prints ONE
JSON line; `--parallelism vocab` runs BASELINE config 4 (vocab rows sharded, NCCL exchange) instead.

`value`  : whole-job tokens/s with inputs resident in HBM.
`e2e`    : the same metric through the host-buffer C entry point (pinned host inputs, H2D and the D2H of
           the metrics inside the timed region, host sync every step like the trainer's `.item()`s).
`roofline`: the lm_head GEMM chain (the dominant kernels), algorithmic 6*T*H*V FLOP per step over the sum
           of its launches' CUDA-event durations, against MEASURED_PEAKS.json.
`--impl reference`: the reference's own algorithm on the host cores (the oracle restatement of its torch
           code; the Python reference itself cannot travel to the GPU box), bounded sample per step.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CONFIGS = {
    # name: (H, V, T, attempts, kl_coef, clamp, batch_size, description)
    "c1": (896, 151936, 8192, 4, 0.0, 10.0, 16, "C1 Qwen2.5-0.5B shape: H=896 V=151936, 16 seqs x 512 packed T=8192"),
    "c2": (1536, 151936, 8192, 16, 0.001, 10.0, 512, "C2 DeepScaleR-1.5B shape: H=1536 V=151936 packed T=8192, 16 attempts/group"),
    "c3": (3584, 152064, 8192, 16, 0.1, 5.0, 4096, "C3 Qwen2.5-7B shape: H=3584 V=152064 packed T=8192, batch 4096 rollouts, PPO + KL to ref"),
    "c4": (5120, 152064, 8192, 16, 0.1, 5.0, 4096, "C4 Qwen2.5-32B shape: H=5120 V=152064 packed T=8192 (vocab-parallel across ranks)"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=32)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS))
    ap.add_argument("--parallelism", default="dp", choices=["dp", "vocab"])
    ap.add_argument("--mode", default="saved", choices=["saved", "recompute"],
                    help="backward: 'saved' = forward keeps bf16 logits (no third GEMM); 'recompute' = nothing [T x V] kept")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-torch-baseline", action="store_true", help="skip the torch-eager-on-this-GPU baseline (N=1 only)")
    ap.add_argument("--cpu-sample-tokens", type=int, default=0, help="tokens per CPU-baseline call (0 = auto)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--exchange", default="symm", choices=["nccl", "symm"],
                    help="vocab-parallel transport: NCCL, or this library's NVSwitch multicast kernels (SymmetricExchange)")
    ap.add_argument("--exchange-ctas", type=int, default=16, help="CTAs of the multicast dH all-reduce kernel (--exchange symm)")
    ap.add_argument("--tunable", action="append", default=[], metavar="NAME=VALUE", help="library tunable (experiments)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# synthetic workload: seeded, packed variable-length sequences, all RL fields
# ------------------------------------------------------------------------------------------------
def packed_lengths(T: int, seed: int) -> list[int]:
    """lognormal(6.2, 0.6) lengths clipped to [64, 4096], greedily packed to exactly T (last one trimmed)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    lens, total = [], 0
    while total < T:
        n = int(np.clip(rng.lognormal(6.2, 0.6), 64, 4096))
        n = min(n, T - total)
        if T - total - n < 8:  # no tiny tail
            n = T - total
        lens.append(n)
        total += n
    return lens


def make_cpu_batch(T: int, V: int, attempts: int, seed: int):
    """CPU tensors [1,T] of every PipelineBatchEncoding field (old/ref logprobs filled in later)."""
    import torch
    from helpers import synthetic_packed_batch
    lens = packed_lengths(T, seed)
    b = synthetic_packed_batch(lens, V, seed=seed, n_prompt=48)
    # group-normalised 0/1-reward advantages per group of `attempts` consecutive sequences
    g = torch.Generator().manual_seed(seed + 1)
    o, k = 0, 0
    while k < len(lens):
        grp = lens[k:k + attempts]
        r = torch.randint(0, 2, (len(grp),), generator=g).float()
        a = (r - r.mean()) / (r.std(unbiased=True).nan_to_num(0) + 1e-4) if len(grp) > 1 else torch.zeros(1)
        for i, n in enumerate(grp):
            b["rewards"][0, o:o + n] = r[i]
            b["advantages"][0, o:o + n] = a[i]
            o += n
        k += attempts
    return b, lens


def reference_step_cpu(hidden, weight, batch, cfg, n_tokens):
    """The reference's algorithm (oracle restatement of rl_step + autograd backward) on the first n_tokens."""
    import torch
    import oracle
    sub = {k: (v[:, :n_tokens].clone() if isinstance(v, torch.Tensor) and v.dim() == 2 else v) for k, v in batch.items()}
    h = hidden[:n_tokens].float().requires_grad_(True)
    w = weight.detach().requires_grad_(True)  # fp32 copy made once by the caller (model weights, not per-step work)
    loss, stats = oracle.oracle_rl_step(h, w, sub, 0, 1, cfg)
    loss.backward()
    return float(loss.detach())


def time_rl_step_api(wl, dev, steps, warmup=3):
    """The drop-in call itself: ``rl_step(model, batch, ...)`` + ``loss.backward()`` through autograd on a model whose
    forward ends in ``lm_head`` (hidden states fixed, as everywhere in this bench), batch already on the device like the
    reference's (finetune_loop.py:109), host sync for the stats dict every step."""
    import types
    import torch
    from pipelinerl_swe_b200 import rl_step
    try:
        class TrunkStub(torch.nn.Module):
            def __init__(self, hidden, weight):
                super().__init__()
                self.hidden = torch.nn.Parameter(hidden[None].clone())
                self.lm_head = torch.nn.Linear(weight.shape[1], weight.shape[0], bias=False, dtype=torch.bfloat16, device=dev)
                self.lm_head.weight.data = weight  # shares the bench's weight storage

            def forward(self, **kw):
                return types.SimpleNamespace(logits=self.lm_head(self.hidden))

        model = TrunkStub(wl.hidden[0], wl.weight)
        batch = wl.batches[0]

        def step():
            loss, stats = rl_step(model, batch, 0, 1, wl.rlcfg)
            loss.backward()
            return stats

        for _ in range(warmup):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            stats = step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out = {"value": wl.T / (ms * 1e-3), "unit": "tokens/s", "ms_per_step": ms, "steps": steps, "loss": stats["loss"],
               "api": "pipelinerl_swe_b200.rl_step(model, batch, step, max_step, config) + loss.backward(); "
                      "bf16 lm_head.weight.grad accumulated by autograd, one host sync per step for the stats dict"}
        # opt-in set_skip_masked_tokens(True): positions whose next token is not labelled stay out of the GEMMs.  Reported
        # apart from every headline number: it does less arithmetic than the metric's 6*T*H*V by the masked fraction.
        from pipelinerl_swe_b200 import set_skip_masked_tokens
        try:
            set_skip_masked_tokens(True)
            for _ in range(warmup):
                step()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(steps):
                step()
            e1.record()
            torch.cuda.synchronize()
            ms2 = e0.elapsed_time(e1) / steps
            lab = batch["labels"][0, 1:] != -100
            out["skip_masked_tokens"] = {"value": wl.T / (ms2 * 1e-3), "unit": "tokens/s", "ms_per_step": ms2,
                                         "labelled_fraction": float(lab.float().mean().item())}
        finally:
            set_skip_masked_tokens(False)
        del model
        torch.cuda.empty_cache()
        return out
    except Exception as e:  # noqa: BLE001
        torch.cuda.empty_cache()
        return {"unavailable": repr(e)[:200]}


def time_torch_eager_gpu(wl, dev, steps=8, warmup=2):
    """The [T, V]-sized part of the reference's rl_step in plain torch on this GPU, bf16 like its HF model:
    lm_head -> /temperature -> log_softmax -> softmax -> entropy -> gather (rl/__init__.py:197-208) and autograd
    backward to hidden and lm_head.weight.  The O(T) per-token loss terms (:213-320) are replaced by one weighted sum,
    so this slightly flatters the reference.  No library of this repo is involved."""
    import torch
    import torch.nn.functional as F
    try:
        hidden = wl.hidden[0].detach().clone().requires_grad_(True)            # bf16 [T-1 or T, H]
        weight = wl.weight.detach().clone().requires_grad_(True)               # bf16 [V, H]
        ids = wl.cpu_batches[0]["input_ids"].to(dev)
        T = ids.shape[1]
        h = None
        tgt = ids[:, 1:].unsqueeze(2)
        c_lp = torch.randn(1, T - 1, device=dev) * 1e-4
        temperature = float(wl.rlcfg.temperature)
        torch.cuda.reset_peak_memory_stats(dev)

        def step():
            logits = F.linear(hidden[None], weight)[:, :-1, :]        # [1, T-1, V]: the model's logits, last position dropped (:200)
            logits = logits / temperature
            logprobs = F.log_softmax(logits, dim=-1)
            probs = F.softmax(logits, dim=-1)
            entropy = -(probs * logprobs).sum(dim=-1)
            del logits, probs
            new_logprobs = torch.gather(logprobs, dim=2, index=tgt).squeeze(2)
            loss = -(new_logprobs.float() * c_lp).sum() - 1e-3 * entropy.float().sum() / (T - 1)
            loss.backward()
            hidden.grad = None
            return loss

        for _ in range(warmup):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out = {"value": T / (ms * 1e-3), "unit": "tokens/s", "ms_per_step": ms, "steps": steps,
               "peak_mem_GB": torch.cuda.max_memory_allocated(dev) / 1e9,
               "kind": "torch eager on the same B200, bf16: lm_head + rl/__init__.py:197-208 + autograd backward "
                       "(weight.grad accumulated); O(T) loss terms omitted"}
        del hidden, weight
        torch.cuda.empty_cache()
        return out
    except Exception as e:  # noqa: BLE001  -- a baseline must never take the bench line down
        torch.cuda.empty_cache()
        return {"unavailable": repr(e)[:200]}


def time_cpu_reference(hidden_cpu, weight_cpu, batch, ocfg, n_tokens, steps, warmup):
    import torch
    torch.set_num_threads(os.cpu_count() or 1)
    weight_cpu = weight_cpu.float()
    for _ in range(warmup):
        reference_step_cpu(hidden_cpu, weight_cpu, batch, ocfg, n_tokens)
    t0 = time.perf_counter()
    for _ in range(steps):
        reference_step_cpu(hidden_cpu, weight_cpu, batch, ocfg, n_tokens)
    dt = (time.perf_counter() - t0) / steps
    return n_tokens / dt, dt, torch.get_num_threads()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines, self.stamped = index, None, [], []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.stamped.append((time.time(), line.strip()))

    def window(self, t0: float, t1: float) -> None:
        """Keep the samples taken inside [t0, t1] (the timed region); fall back to everything seen under load."""
        inside = [ln for t, ln in self.stamped if t0 <= t <= t1]
        self.lines = inside or [ln for _, ln in self.stamped[-5:]]

    def __exit__(self, *exc):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        return False

    def summary(self) -> dict:
        sm, mx, pw, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw), "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# the two arms
# ------------------------------------------------------------------------------------------------
def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"burst": d["bf16_tflops"], "sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "hbm": d["hbm_gbs"],
                "source": "measured"}
    return {"burst": 1590.0, "sustained": 1400.0, "hbm": 6650.0, "source": "fallback"}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port) on the host cores."""
    import torch
    rank, world, _ = dist_env()
    if rank != 0:
        return
    import oracle
    H, V, T, attempts, kl, clamp, B, desc = CONFIGS[args.config]
    n_tok = args.cpu_sample_tokens or max(64, min(512, (1 << 31) // (V * 4 * 6)))
    batch, _ = make_cpu_batch(T, V, attempts, seed=1236)
    g = torch.Generator().manual_seed(1236)
    hidden = torch.randn(n_tok, H, generator=g).bfloat16()
    weight = (torch.randn(V, H, generator=g) * 0.02).bfloat16()
    ocfg = oracle.OracleRLConfig(policy_loss="ppo", batch_size=B, kl_coef=kl, final_kl_coef=kl, clamp_log_ratio_ref_new_value=clamp)
    tps, dt, threads = time_cpu_reference(hidden, weight, batch, ocfg, n_tok, args.steps, args.warmup)
    sample = f"first {n_tok} tokens of the {args.config} packed row per step (fp32 math on bf16 values, fwd + autograd bwd)"
    line = {
        "impl": "reference", "metric": "grpo_logprob_loss_fwd_bwd_tokens_per_sec", "value": tps, "unit": "tokens/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "H": H, "V": V, "T": T, "tokens_per_step_sample": n_tok},
        "cpu_baseline": {"value": tps, "unit": "tokens/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": tps, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


class DeviceWorkload:
    """Everything one rank needs resident in HBM for the timed region (and pinned host copies for e2e)."""

    def __init__(self, args, dev, rank, world):
        import torch
        from pipelinerl_swe_b200 import RLConfig, ops
        from pipelinerl_swe_b200 import rl as prl
        from pipelinerl_swe_b200.vocab_parallel import shard_range
        self.torch, self.ops, self.prl = torch, ops, prl
        H, V, T, attempts, kl, clamp, B, desc = CONFIGS[args.config]
        self.H, self.V, self.T, self.desc = H, V, T, desc
        self.vocab_parallel = args.parallelism == "vocab" and world > 1
        self.exchange = None
        if self.vocab_parallel and args.exchange == "symm":
            from pipelinerl_swe_b200.vocab_parallel import SymmetricExchange
            try:
                self.exchange = SymmetricExchange(None, dev, T, H, allreduce_ctas=args.exchange_ctas)
            except RuntimeError as e:   # no NVSwitch multicast mapping on this box: NCCL transport, reported in `config.exchange`
                print(f"bench.py: {e}", file=sys.stderr)
        seed = 1236 + (0 if self.vocab_parallel else rank)
        self.rlcfg = RLConfig(policy_loss="ppo", batch_size=B, kl_coef=kl, final_kl_coef=kl, clamp_log_ratio_ref_new_value=clamp,
                              epsilon=0.2, temperature=1.0)
        self.cfg = prl.make_loss_cfg(self.rlcfg, 0, 1, True, False)
        g = torch.Generator(device=dev).manual_seed(seed)
        self.n_batches = 2
        self.hidden = [torch.randn(T, H, device=dev, generator=g).bfloat16() for _ in range(self.n_batches)]
        gw = torch.Generator(device=dev).manual_seed(99)  # same weights on every rank
        wfull_rows = V
        self.v_offset = 0
        if self.vocab_parallel:
            b, e = shard_range(V, world, rank)
            self.v_offset, wfull_rows = b, e - b
            full = (torch.randn(V, H, device=dev, generator=gw) * 0.02).bfloat16()
            self.weight = full[b:e].contiguous()
            del full
        else:
            self.weight = (torch.randn(V, H, device=dev, generator=gw) * 0.02).bfloat16()
        self.Vloc = wfull_rows
        self.grad_weight = torch.zeros(self.Vloc, H, dtype=torch.float32, device=dev)
        self.cpu_batches, self.batches, self.targets = [], [], []
        for i in range(self.n_batches):
            cb, lens = make_cpu_batch(T, V, attempts, seed=seed * 10 + i)
            db = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in cb.items()}
            tgt = db["input_ids"][0, 1:].contiguous()
            # old/ref logprobs near the current policy: needs the policy's logprobs once
            stats = self.fwd_stats(self.hidden[i][:-1], tgt)
            lp, ent, lse = ops.lmhead_finalize(stats)
            m = (db["labels"][0, 1:] != -100).float()
            gn = torch.Generator(device=dev).manual_seed(seed * 10 + i)
            db["old_logprobs"][0, 1:] = (lp + 0.05 * torch.randn(T - 1, device=dev, generator=gn)) * m
            db["ref_logprobs"][0, 1:] = (lp + 0.10 * torch.randn(T - 1, device=dev, generator=gn)) * m
            for k in ("old_logprobs", "ref_logprobs"):
                cb[k] = db[k].cpu()
            self.cpu_batches.append(cb); self.batches.append(db); self.targets.append(tgt)
        self.hidden_m = [h[:-1] for h in self.hidden]  # rows that have a next token (contiguous view)
        self.save_logits = args.mode != "recompute"  # default: forward keeps bf16 logits, backward skips the recompute GEMM
        self.saved = ops.lmhead_saved_buffer(T - 1, H, self.Vloc, dev) if self.save_logits else None
        torch.cuda.synchronize()

    def fwd_stats(self, h, tgt, saved=None):
        if self.exchange is not None:   # the merge kernel's multimem.st is the all-gather
            return self.exchange.forward_stats(
                lambda mc: self.ops.lmhead_fwd_stats(h, self.weight, tgt, 1.0, self.v_offset, saved, stats_mc=mc), h.shape[0])
        stats = self.ops.lmhead_fwd_stats(h, self.weight, tgt, 1.0, self.v_offset, saved)
        if self.vocab_parallel:
            from pipelinerl_swe_b200.vocab_parallel import exchange_stats
            stats = exchange_stats(stats)
        return stats

    def step(self, i: int):
        """One device-resident fwd+bwd micro-batch; returns the device metrics vector."""
        ops, torch = self.ops, self.torch
        k = i % self.n_batches
        h, tgt, b = self.hidden_m[k], self.targets[k], self.batches[k]
        saved = self.saved if self.save_logits else None
        lp, ent, lse = ops.lmhead_finalize(self.fwd_stats(h, tgt, saved))
        loss, metrics, coef_lp, coef_ent, _ = ops.token_loss(lp[None], ent[None], b, self.cfg)
        args = (h, self.weight, tgt, lse, ent, coef_lp[0], coef_ent[0], 1.0, self.v_offset)
        if self.exchange is not None:   # dH partial into the symmetric buffer, in-switch all-reduce beside the dW GEMM
            from pipelinerl_swe_b200 import _abi
            x, M = self.exchange, h.shape[0]
            ops.lmhead_bwd(*args, need_grad_weight=False, saved=saved, stages=_abi.BWD_DLOGITS | _abi.BWD_DH, grad_hidden_out=x.dh_view(M))
            done = x.allreduce_dh_async(M)
            ops.lmhead_bwd(*args, need_grad_hidden=False, grad_weight_out=self.grad_weight, accumulate=True, saved=saved,
                           stages=_abi.BWD_DW)
            torch.cuda.current_stream().wait_event(done)
        elif self.vocab_parallel:   # as LMHeadLogprobFunction.backward: the dH all-reduce runs while dW is computed
            from pipelinerl_swe_b200 import _abi
            from pipelinerl_swe_b200.vocab_parallel import reduce_grad_hidden_async
            gh, _ = ops.lmhead_bwd(*args, need_grad_weight=False, saved=saved, stages=_abi.BWD_DLOGITS | _abi.BWD_DH)
            pending = reduce_grad_hidden_async(gh)
            ops.lmhead_bwd(*args, need_grad_hidden=False, grad_weight_out=self.grad_weight, accumulate=True, saved=saved,
                           stages=_abi.BWD_DW)
            if pending is not None:
                pending.wait()
        else:
            ops.lmhead_bwd(*args, grad_weight_out=self.grad_weight, accumulate=True, saved=saved)
        return metrics


def run_b200(args):
    import torch
    import torch.distributed as dist
    from pipelinerl_swe_b200 import _abi
    from pipelinerl_swe_b200.host_step import HostRLStep

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- pipelinerl_swe_b200 has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")  # keep rank 0's stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    _abi.set_tunable("save_logits", 0 if args.mode == "recompute" else 1)
    for kv in args.tunable:
        name, val = kv.split("=")
        _abi.set_tunable(name, int(val))

    wl = DeviceWorkload(args, dev, rank, world)
    H, V, T = wl.H, wl.V, wl.T
    K, W = args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident value ----------------
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:  # started before the warm-up: nvidia-smi needs a moment to come up
        for i in range(W):
            wl.step(i)
        barrier()
        _abi.profile_enable(16 * K + 64)
        barrier()
        t_begin = time.time()
        n_launch0 = _abi.launch_count()
        e0.record()
        for i in range(K):
            metrics = wl.step(i)
        e1.record()
        n_launch = _abi.launch_count() - n_launch0   # counted by the library at every kernel launch it issues
        barrier()
        t_end = time.time()
        prof = _abi.profile_read(16 * K + 64)
        _abi.profile_enable(0)
        ms_total = max_over_ranks(e0.elapsed_time(e1))   # identical on every rank: the branch below stays collective-safe
        if ms_total < 350.0:         # very short timed regions: give the 100 ms sampler a chance under the same load
            for i in range(int(400.0 / max(ms_total / K, 1e-3)) + 1):   # same count on every rank (ms_total is all-reduced)
                wl.step(i)
            barrier()
            t_end = time.time()
    clocks.window(t_begin, t_end)
    m = metrics.tolist()
    assert m[_abi.M_FLAGS] == 0 and m[_abi.M_N_MASKED] > 0, "bench workload produced non-finite values"
    ms_step = ms_total / K
    tokens_per_step = T * (1 if wl.vocab_parallel else world)
    value = tokens_per_step * K / (ms_total * 1e-3)

    per_tag: dict[str, float] = {}
    for tag, ms in prof:
        per_tag[tag] = per_tag.get(tag, 0.0) + ms / K
    gemm_tags = ("fwd_gemm_stats", "bwd_recompute_dlogits", "bwd_dH_gemm", "bwd_dW_gemm")
    gemm_only_ms = sum(per_tag.get(t, 0.0) for t in gemm_tags)
    gemm_ms = sum(per_tag.values())  # the whole chain: GEMMs + merge + finalize + token loss + dlogits pass + split-K reduce
    n_gemms = sum(1 for t in gemm_tags if t in per_tag)
    peaks = load_peaks()
    algo_flops = 6.0 * (T - 1) * H * wl.Vloc
    achieved = algo_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    executed = 2.0 * (T - 1) * H * wl.Vloc * n_gemms / (gemm_only_ms * 1e-3) / 1e12 if gemm_only_ms > 0 else 0.0
    traffic, traffic_note = None, "no ncu capture for this workload"
    try:  # DRAM bytes of the step's kernels from the committed ncu capture (C3, saved mode, unsharded only)
        tj = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01", "traffic.json")))
        if tj["workload"] == f"{args.config}/{args.mode}/{args.parallelism}":
            traffic = float(sum(tj["dram_bytes_per_launch"].values()))
            traffic_note = ("dram__bytes_read+write summed over the step's four large kernels (one launch each per step), " + tj["source"]
                            + "; algorithmic operand+output bytes of the same kernels: %.1f GB" % (sum(tj["algorithmic_bytes_per_launch"].values()) / 1e9))
    except Exception:  # noqa: BLE001
        pass
    roofline = {
        "bound": "tensor", "achieved": achieved, "peak": peaks["sustained"], "unit": "TFLOP/s", "frac": achieved / peaks["sustained"],
        "traffic": traffic, "traffic_note": traffic_note, "peak_kind": f"bf16_tflops_sustained of {peaks['source']} (kernels timed inside a long step)",
        "frac_of_burst_peak": achieved / peaks["burst"], "executed_tflops": executed, "gemm_ms_per_step": gemm_ms,
        "share_of_step": gemm_ms / ms_step if world == 1 else None,
        "kernels_ms_per_step": {k: round(v, 4) for k, v in sorted(per_tag.items())},
        "gemm_only_ms_per_step": gemm_only_ms,
        "note": "achieved = algorithmic 6*T*H*V FLOP per step / summed CUDA-event time of EVERY kernel of the step "
                "(tcgen05 GEMMs + merge + finalize + token loss + dlogits pass + split-K reduce)",
    }
    launches_per_step = n_launch / K if K else 0

    # ---------------- end to end through the host-buffer C entry point ----------------
    e2e = None
    if not args.no_e2e and not wl.vocab_parallel:
        hs = [HostRLStep(T, H, V, wl.weight, wl.rlcfg, 0, 1, grad_weight_dtype=torch.float32) for _ in range(2)]
        for k, s in enumerate(hs):
            s.load_batch(wl.cpu_batches[k % wl.n_batches], wl.hidden[k % wl.n_batches].cpu())
        # Two streams, two pinned buffer sets: step i+1 is submitted (its H2D copies included) before the host
        # waits for step i's metrics -- the role of the reference's loader thread, which moves the next batch to the
        # device while the current one trains (finetune_loop.py:92-115).  Every step still copies its own inputs
        # from pinned memory and the host still reads every step's metrics before it looks at the next one.
        streams = [torch.cuda.Stream(device=dev) for _ in range(2)]

        def submit(i):
            with torch.cuda.stream(streams[i % 2]):
                return hs[i % 2].run(accumulate=True)

        for i in range(W):
            submit(i)
            streams[i % 2].synchronize()
        barrier()
        cur = torch.cuda.current_stream()
        e0.record(cur)
        for st in streams:
            st.wait_stream(cur)
        mh = submit(0)
        for i in range(K):
            mh_next = submit(i + 1) if i + 1 < K else None
            streams[i % 2].synchronize()               # the trainer reads the step's metrics on the host every step
            _ = float(mh[0])
            mh = mh_next
        for st in streams:
            cur.wait_stream(st)
        e1.record(cur)
        barrier()
        ms_e2e = max_over_ranks(e0.elapsed_time(e1))
        e2e = {"value": T * world * K / (ms_e2e * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": hs[0].h2d_bytes,
               "d2h_bytes_per_step": hs[0].d2h_bytes, "ms_per_step": ms_e2e / K,
               "api": "b200grpo_rl_step_host (pinned host hidden + batch fields in, f64 metrics out; weight/grad stay on device); "
                      "two streams: step i+1 is submitted before the host waits for step i's metrics"}
        del hs
        torch.cuda.empty_cache()

    # ---------------- the reference's torch ops on the SAME GPU (rank 0, N=1 only; SURVEY 8d "like-for-like") ----------------
    torch_gpu, api = None, None
    if rank == 0 and world == 1 and not args.no_torch_baseline:
        api = time_rl_step_api(wl, dev, K)
        torch_gpu = time_torch_eager_gpu(wl, dev)

    # ---------------- CPU baseline (rank 0, N=1 only) ----------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        import oracle
        n_tok = args.cpu_sample_tokens or 1024
        ocfg = oracle.OracleRLConfig(**wl.rlcfg.model_dump())
        tps, dt, threads = time_cpu_reference(wl.hidden[0][:n_tok].cpu(), wl.weight.cpu(), wl.cpu_batches[0], ocfg, n_tok, 3, 1)
        cpu = {"value": tps, "unit": "tokens/s", "cores": threads, "kind": "port",
               "sample": f"first {n_tok} tokens of the same packed row, 1 warm-up + 3 timed fwd+bwd calls ({dt:.2f} s each), fp32 torch math on the bf16 values"}

    if rank == 0:
        line = {
            "metric": "grpo_logprob_loss_fwd_bwd_tokens_per_sec", "value": value, "unit": "tokens/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if wl.vocab_parallel else "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": wl.desc, "H": H, "V": V, "T": T, "parallelism": ("vocab" if wl.vocab_parallel else "dp") + str(world),
                       **({"exchange": "nvswitch multicast kernels (SymmetricExchange)" if wl.exchange is not None else "nccl"} if wl.vocab_parallel else {}),
                       "grad_weight": "f32 accumulated across steps", "l2": "inputs exceed L2 (lm_head weight %.2f GB)" % (wl.weight.numel() * 2 / 1e9),
                       "bwd_mode": args.mode, "tunables": args.tunable, "tokens_per_gpu_per_sec": value / world if not wl.vocab_parallel else value},
            "roofline": roofline, "cpu_baseline": cpu, "torch_gpu_baseline": torch_gpu, "rl_step_api": api, "e2e": e2e, "gpu_launches": int(round(launches_per_step * K)),
            "gpu_launches_per_step": launches_per_step, "clocks": clocks.summary(),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
