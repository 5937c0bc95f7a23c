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
`parity`  : before anything is timed, the exact kernels of the step are compared at the benchmarked shape with the
           reference's fp32 math (tests/shape_parity.py): every row's logprob/entropy, the full dH and the full dW.
`recompute`: the same workload with `--mode recompute` (nothing [T x V] kept between forward and backward).
`vocab_parallel` (N>1): BASELINE config 4 (H=5120) with the vocab rows sharded over the N ranks, both transports,
           parity-checked against the unsharded fp32 reference, with the unsharded 1-GPU time beside it.
`--impl reference`: the reference's OWN rl_step (oracle/_ref, staged by oracle/build_ref.py) + backward() on the host
           cores, bounded sample per step; falls back to the oracle restatement only if oracle/_ref is absent.
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
    ap.add_argument("--steps", type=int, default=96)
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
    ap.add_argument("--no-parity", action="store_true", help="skip the pre-run parity gate at the benchmarked shape")
    ap.add_argument("--no-recompute", action="store_true", help="skip the recompute-mode sub-record")
    ap.add_argument("--no-vocab-record", action="store_true", help="N>1: skip the config-4 vocab-parallel sub-record")
    ap.add_argument("--no-weight-record", action="store_true", help="N>1: skip the config-5 weight-broadcast sub-record")
    ap.add_argument("--no-other-configs", action="store_true", help="N=1: skip the short C1 / C2 / C4-unsharded records")
    ap.add_argument("--cpu-budget-s", type=float, default=0.0, help="wall-clock budget of the CPU arm (0 = auto)")
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


def rl_config_kwargs(name: str) -> dict:
    H, V, T, attempts, kl, clamp, B, desc = CONFIGS[name]
    return dict(policy_loss="ppo", batch_size=B, kl_coef=kl, final_kl_coef=kl, clamp_log_ratio_ref_new_value=clamp, epsilon=0.2,
                temperature=1.0)


class CpuArm:
    """The reference's CPU implementation of the path on the host cores: its own ``rl_step`` (oracle/_ref, the files of
    /root/reference/pipelinerl/finetune/rl staged by oracle/build_ref.py) + ``loss.backward()`` on a model that is only
    the lm_head (fp32 math on the bf16 values, gradients accumulated in ``.grad`` like the trainer's micro-batches).
    Falls back to the oracle restatement (kind "port") only when oracle/_ref is not there, and says why."""

    def __init__(self, name: str, hidden, weight, batch):
        import torch
        self.torch = torch
        torch.set_num_threads(os.cpu_count() or 1)
        self.threads = torch.get_num_threads()
        self.hidden, self.batch = hidden.float(), batch
        self.weight = torch.nn.Parameter(weight.float())   # fp32 copy made once (model state, not per-step work)
        kw = rl_config_kwargs(name)
        try:
            from oracle import build_ref
            self.rl = build_ref.load()
            self.PBE = sys.modules["pipelinerl.finetune.types"].PipelineBatchEncoding
            self.cfg = self.rl.RLConfig(**kw)
            self.kind, self.why = "reference", "pipelinerl.finetune.rl.rl_step from oracle/_ref (unmodified reference files)"
        except Exception as e:  # noqa: BLE001
            import oracle
            self.oracle, self.cfg = oracle, oracle.OracleRLConfig(**kw)
            self.kind, self.why = "port", f"oracle restatement (oracle/_ref unavailable: {e!r})"[:200]
            print(f"bench.py: CPU arm falls back to the oracle port: {e!r}", file=sys.stderr)

    def step(self, n_tokens: int) -> float:
        torch = self.torch
        sub = {k: (v[:, :n_tokens].clone() if isinstance(v, torch.Tensor) and v.dim() == 2 else v) for k, v in self.batch.items()}
        h = self.hidden[:n_tokens].clone().requires_grad_(True)
        if self.kind == "reference":
            import types
            w = self.weight

            class LMHeadOnly(torch.nn.Module):
                def forward(self, **kw):
                    return types.SimpleNamespace(logits=torch.nn.functional.linear(h[None], w))

            sub.pop("seq_boundaries", None)
            loss, _ = self.rl.rl_step(LMHeadOnly(), self.PBE(**sub, model_version=0), 0, 1, self.cfg)
        else:
            loss, _ = self.oracle.oracle_rl_step(h, self.weight, sub, 0, 1, self.cfg)
        loss.backward()
        return float(loss.detach())

    def run(self, T: int, steps: int, warmup: int, budget_s: float, fixed_tokens: int = 0):
        """Pick the largest sample (multiple of 1024 tokens, up to the whole row) whose warmup + steps fit the budget."""
        t0 = time.perf_counter()
        self.step(min(1024, T))
        probe = time.perf_counter() - t0          # includes first-touch of the fp32 gradient: an over-estimate
        n_tok = fixed_tokens or int(max(1, min(T // 1024, budget_s / ((steps + warmup) * probe)))) * 1024
        n_tok = min(n_tok, T)
        for _ in range(warmup):
            self.step(n_tok)
        t0 = time.perf_counter()
        for _ in range(steps):
            loss = self.step(n_tok)
        dt = (time.perf_counter() - t0) / steps
        whole = n_tok >= T
        sample = (f"{'the whole' if whole else 'first ' + str(n_tok) + ' tokens of the'} {T}-token packed row per step, {warmup} warm-up + "
                  f"{steps} timed rl_step + backward() calls ({dt:.2f} s each), fp32 torch math on the bf16 values, {self.why}")
        return {"value": n_tok / dt, "unit": "tokens/s", "cores": self.threads, "kind": self.kind, "sample": sample,
                "tokens_per_step_sample": n_tok, "same_config": bool(whole), "s_per_step": dt, "loss": loss}


def cpu_workload(name: str, n_tokens: int):
    """The CPU arm's own copy of the workload when no GPU is involved (--impl reference): same generator as the GPU arm's
    batch 0, CPU-generated hidden states, old/ref log-probs near the current policy."""
    import torch
    H, V, T, attempts, kl, clamp, B, desc = CONFIGS[name]
    batch, _ = make_cpu_batch(T, V, attempts, seed=12360)
    g = torch.Generator().manual_seed(1236)
    hidden = torch.randn(T, H, generator=g).bfloat16()
    weight = (torch.randn(V, H, generator=g) * 0.02).bfloat16()
    n = min(n_tokens, T)
    with torch.no_grad():   # the policy's own log-probs of the first n tokens (one chunked fp32 forward, untimed)
        wf = weight.float()
        lp = torch.zeros(T - 1)
        for r0 in range(0, n - 1, 512):
            r1 = min(n - 1, r0 + 512)
            z = hidden[r0:r1].float() @ wf.t()
            lp[r0:r1] = torch.log_softmax(z, -1).gather(1, batch["input_ids"][0, r0 + 1:r1 + 1, None]).squeeze(1)
        m = (batch["labels"][0, 1:] != -100).float()
        batch["old_logprobs"][0, 1:] = (lp + 0.05 * torch.randn(T - 1, generator=g)) * m
        batch["ref_logprobs"][0, 1:] = (lp + 0.10 * torch.randn(T - 1, generator=g)) * m
    return hidden, weight, batch


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
        from pipelinerl_swe_b200 import set_fused_grad_accumulation, set_save_logits
        set_save_logits(wl.mode != "recompute")

        def step():
            loss, stats = rl_step(model, batch, 0, 1, wl.rlcfg)
            loss.backward()
            return stats

        def timed():
            for _ in range(warmup):
                step()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                stats = step()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / steps, stats

        # headline of this record: the opt-in fused f32 accumulation (dW added straight into lm_head.weight.main_grad by the
        # GEMM epilogue, Megatron main_grad convention); then the default, where autograd adds a fresh bf16 dW into .grad
        set_fused_grad_accumulation(True)
        try:
            ms, stats = timed()
        finally:
            set_fused_grad_accumulation(False)
        model.lm_head.weight.main_grad = None
        torch.cuda.empty_cache()
        ms_bf16, _ = timed()
        out = {"value": wl.T / (ms * 1e-3), "unit": "tokens/s", "ms_per_step": ms, "steps": steps, "loss": stats["loss"],
               "api": "pipelinerl_swe_b200.rl_step(model, batch, step, max_step, config) + loss.backward() with "
                      "set_fused_grad_accumulation(True): f32 lm_head.weight.main_grad accumulated by the dW GEMM epilogue, "
                      "one host sync per step for the stats dict",
               "default_bf16_autograd_grad": {"value": wl.T / (ms_bf16 * 1e-3), "unit": "tokens/s", "ms_per_step": ms_bf16,
                                              "note": "set_fused_grad_accumulation(False): backward returns a fresh bf16 [V,H] dW, autograd adds it to .grad"}}
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
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


def time_cublas_same_box(dev, secs=2.0):
    """The peak's own recipe on THIS box in THIS run: torch.matmul bf16 8192^3 back to back (MEASURED_PEAKS.json `how`: sustained).
    Boxes of the pool differ by up to 13 % under the 1 kW cap; this lets `roofline.frac` be read against the box that ran the step.
    A library GEMM as a yardstick only -- nothing of the measured path touches it."""
    import torch
    try:
        n = 8192
        a = torch.randn(n, n, device=dev, dtype=torch.bfloat16)
        b = torch.randn(n, n, device=dev, dtype=torch.bfloat16)
        c = torch.empty(n, n, device=dev, dtype=torch.bfloat16)
        for _ in range(20):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 0
        t0 = time.time()
        e0.record()
        while time.time() - t0 < secs:
            for _ in range(50):
                torch.matmul(a, b, out=c)
            iters += 50
            torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        del a, b, c
        torch.cuda.empty_cache()
        return {"bf16_tflops_sustained": 2.0 * n ** 3 / (ms * 1e-3) / 1e12, "seconds": secs, "how": "torch.matmul bf16 8192^3 back to back, this box, this run"}
    except Exception as e:  # noqa: BLE001
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
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref) on the host cores."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    H, V, T, attempts, kl, clamp, B, desc = CONFIGS[args.config]
    K, W = args.steps, args.warmup
    budget = args.cpu_budget_s or 240.0
    hidden, weight, batch = cpu_workload(args.config, T)
    arm = CpuArm(args.config, hidden, weight, batch)
    r = arm.run(T, K, W, budget, args.cpu_sample_tokens)
    line = {
        "impl": "reference", "metric": "grpo_logprob_loss_fwd_bwd_tokens_per_sec", "value": r["value"], "unit": "tokens/s",
        "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": r["s_per_step"] * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "H": H, "V": V, "T": T, "tokens_per_step_sample": r["tokens_per_step_sample"],
                   "same_config": r["same_config"]},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


class DeviceWorkload:
    """Everything one rank needs resident in HBM for the timed region (and pinned host copies for e2e)."""

    def __init__(self, name, dev, rank, world, *, vocab_parallel=False, exchange="symm", exchange_ctas=16, mode="saved",
                 keep_full=False):
        import torch
        from pipelinerl_swe_b200 import RLConfig, ops
        from pipelinerl_swe_b200 import rl as prl
        from pipelinerl_swe_b200.vocab_parallel import shard_range
        self.torch, self.ops, self.prl = torch, ops, prl
        H, V, T, attempts, kl, clamp, B, desc = CONFIGS[name]
        self.name, self.H, self.V, self.T, self.desc = name, H, V, T, desc
        self.rank, self.world = rank, world
        self.vocab_parallel = vocab_parallel and world > 1
        self.exchange = None
        if self.vocab_parallel and exchange == "symm":
            from pipelinerl_swe_b200.vocab_parallel import SymmetricExchange
            try:
                self.exchange = SymmetricExchange(None, dev, T, H, allreduce_ctas=exchange_ctas)
            except RuntimeError as e:   # no NVSwitch multicast mapping on this box: NCCL transport, reported in `config.exchange`
                print(f"bench.py: {e}", file=sys.stderr)
        seed = 1236 + (0 if self.vocab_parallel else rank)
        self.rlcfg = RLConfig(**rl_config_kwargs(name))
        self.cfg = prl.make_loss_cfg(self.rlcfg, 0, 1, True, False)
        g = torch.Generator(device=dev).manual_seed(seed)
        self.n_batches = 2
        self.hidden = [torch.randn(T, H, device=dev, generator=g).bfloat16() for _ in range(self.n_batches)]
        gw = torch.Generator(device=dev).manual_seed(99)  # same weights on every rank
        self.v_offset, self.Vloc, self.weight_full = 0, V, None
        full = (torch.randn(V, H, device=dev, generator=gw) * 0.02).bfloat16()
        if self.vocab_parallel:
            b, e = shard_range(V, world, rank)
            self.v_offset, self.Vloc = b, e - b
            self.weight = full[b:e].contiguous()
            if keep_full:
                self.weight_full = full   # only for the parity gate (the reference needs every vocab row)
        else:
            self.weight = self.weight_full = full
        del full
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
        self.set_mode(mode)
        torch.cuda.synchronize()

    def set_mode(self, mode: str) -> None:
        """'saved': forward keeps bf16 logits, backward skips the recompute GEMM; 'recompute': nothing [T x V] is kept."""
        self.mode = mode
        self.save_logits = mode != "recompute"
        self.saved = self.ops.lmhead_saved_buffer(self.T - 1, self.H, self.Vloc, self.hidden[0].device) if self.save_logits else None

    def fwd_stats(self, h, tgt, saved=None):
        if self.exchange is not None:   # the merge kernel's multimem.st is the all-gather
            return self.exchange.forward_stats(
                lambda mc: self.ops.lmhead_fwd_stats(h, self.weight, tgt, 1.0, self.v_offset, saved, stats_mc=mc), h.shape[0])
        stats = self.ops.lmhead_fwd_stats(h, self.weight, tgt, 1.0, self.v_offset, saved)
        if self.vocab_parallel:
            from pipelinerl_swe_b200.vocab_parallel import exchange_stats
            stats = exchange_stats(stats)
        return stats

    def step(self, i: int, detail: bool = False):
        """One device-resident fwd+bwd micro-batch; returns the device metrics vector (detail: everything the gate compares)."""
        ops, torch = self.ops, self.torch
        k = i % self.n_batches
        h, tgt, b = self.hidden_m[k], self.targets[k], self.batches[k]
        saved = self.saved if self.save_logits else None
        lp, ent, lse = ops.lmhead_finalize(self.fwd_stats(h, tgt, saved))
        loss, metrics, coef_lp, coef_ent, _ = ops.token_loss(lp[None], ent[None], b, self.cfg)
        # no entropy bonus in this config: K3's dLoss/dentropy is identically zero and the backward is told so (as rl_step does)
        c_ent = coef_ent[0] if self.cfg.entropy_coef != 0.0 else None
        args = (h, self.weight, tgt, lse, ent, coef_lp[0], c_ent, 1.0, self.v_offset)
        if self.exchange is not None:   # dH partial into the symmetric buffer, in-switch all-reduce beside the dW GEMM
            from pipelinerl_swe_b200 import _abi
            x, M = self.exchange, h.shape[0]
            gh = x.dh_view(M)
            ops.lmhead_bwd(*args, need_grad_weight=False, saved=saved, stages=_abi.BWD_DLOGITS | _abi.BWD_DH, grad_hidden_out=gh)
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
            gh, _ = ops.lmhead_bwd(*args, grad_weight_out=self.grad_weight, accumulate=True, saved=saved)
        if detail:
            return dict(metrics=metrics, lp=lp, ent=ent, lse=lse, coef_lp=coef_lp[0], coef_ent=coef_ent[0], gh=gh, loss=loss)
        return metrics

    def parity(self) -> dict:
        """The pre-run gate: ONE step of exactly what is timed (this object's `step`, default tunables, this mode, this
        transport) into a zeroed f32 gradient, then a second one on top of it, against the reference's fp32 math on the
        same bf16 values (tests/shape_parity.py: rl/__init__.py:197-210 + autograd, full vocab, row-chunked)."""
        import shape_parity as sp
        torch = self.torch
        import torch.distributed as dist
        self.grad_weight.zero_()
        d = self.step(0, detail=True)
        self.step(0)                                     # accumulate=True on a non-zero gradient
        gw = self.grad_weight / 2
        h, tgt = self.hidden_m[0], self.targets[0]
        ref = sp.reference_lmhead(h, self.weight_full, tgt, 1.0, d["coef_lp"], d["coef_ent"])
        lo, hi = self.v_offset, self.v_offset + self.Vloc
        rec = {"lp_rel": sp.rel_max(d["lp"], ref["lp"]), "ent_rel": sp.rel_max(d["ent"], ref["ent"]),
               "dH_rel": sp.rel_max(d["gh"], ref["dH"]), "dW_rel": sp.rel_max(gw, ref["dW"][lo:hi]),
               "fast_reference_vs_plain_fp32": sp.fast_reference_is_fp32(h, self.weight_full, tgt, 1.0),
               "mode": self.mode, "rows": int(h.shape[0]), "accumulate_steps": 2,
               "against": "fp32 log_softmax/entropy/gather + autograd of rl/__init__.py:197-210 on the same bf16 values; every row, full dH, full dW"}
        tr = sp.tail_tile_rows(self.Vloc, self.H, 256)
        if tr is not None:
            rec["dW_tail_split_tiles_rel"] = float((gw[tr[0]:tr[1]] - ref["dW"][lo + tr[0]:lo + tr[1]]).abs().max() / ref["dW"].abs().max())
        ok = rec["lp_rel"] < sp.LP_RTOL and rec["ent_rel"] < sp.LP_RTOL and rec["dH_rel"] < sp.GRAD_RTOL and rec["dW_rel"] < sp.GRAD_RTOL
        if self.vocab_parallel:   # every rank must hold bit-identical log-probs (K3 then needs no exchange) and pass
            mx, mn = d["lp"].clone(), d["lp"].clone()
            dist.all_reduce(mx, op=dist.ReduceOp.MAX); dist.all_reduce(mn, op=dist.ReduceOp.MIN)
            rec["lp_identical_across_ranks"] = bool(torch.equal(mx, mn))
            ok = ok and rec["lp_identical_across_ranks"]
            t = torch.tensor([rec["lp_rel"], rec["ent_rel"], rec["dH_rel"], rec["dW_rel"], 0.0 if ok else 1.0], dtype=torch.float64,
                             device=h.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)     # worst rank
            rec.update(lp_rel=float(t[0]), ent_rel=float(t[1]), dH_rel=float(t[2]), dW_rel=float(t[3]))
            ok = float(t[4]) == 0.0
        rec["ok"] = bool(ok)
        self.grad_weight.zero_()
        del ref
        torch.cuda.empty_cache()
        return rec


class Timer:
    """Barrier + sync on both sides, CUDA events on the launching stream, max over ranks."""

    def __init__(self, dev, world):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.dev, self.world = torch, dist, dev, world

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def time_steps(self, wl, K: int, W: int, profile: bool = True):
        """W untimed + exactly K timed steps of ``wl``.  Returns (ms_total, per-kernel ms/step, launches, last metrics, t0, t1)."""
        from pipelinerl_swe_b200 import _abi
        torch = self.torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for i in range(W):
            wl.step(i)
        self.barrier()
        if profile:
            _abi.profile_enable(16 * K + 64)
        self.barrier()
        t_begin = time.time()
        n0 = _abi.launch_count()
        e0.record()
        for i in range(K):
            metrics = wl.step(i)
        e1.record()
        n_launch = _abi.launch_count() - n0   # counted by the library at every kernel launch it issues
        self.barrier()
        t_end = time.time()
        per_tag: dict[str, float] = {}
        if profile:
            for tag, ms in _abi.profile_read(16 * K + 64):
                per_tag[tag] = per_tag.get(tag, 0.0) + ms / K
            _abi.profile_enable(0)
        ms_total = self.max_over_ranks(e0.elapsed_time(e1))
        return ms_total, per_tag, n_launch, metrics, t_begin, t_end


GEMM_TAGS = ("fwd_gemm_stats", "bwd_recompute_dlogits", "bwd_dH_gemm", "bwd_dW_gemm")


def roofline_record(per_tag, ms_step, T, H, Vloc, peaks, world):
    gemm_only_ms = sum(per_tag.get(t, 0.0) for t in GEMM_TAGS)
    gemm_ms = sum(per_tag.values())  # the whole chain: GEMMs + merge + finalize + token loss + dlogits pass + tail reduce
    n_gemms = sum(1 for t in GEMM_TAGS if t in per_tag)
    algo_flops = 6.0 * (T - 1) * H * Vloc
    achieved = algo_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    executed = 2.0 * (T - 1) * H * Vloc * n_gemms / (gemm_only_ms * 1e-3) / 1e12 if gemm_only_ms > 0 else 0.0
    return {
        "bound": "tensor", "achieved": achieved, "peak": peaks["sustained"], "unit": "TFLOP/s", "frac": achieved / peaks["sustained"],
        "peak_kind": f"bf16_tflops_sustained of {peaks['source']} (kernels timed inside a long step)",
        "frac_of_burst_peak": achieved / peaks["burst"], "burst_peak": peaks["burst"], "executed_tflops": executed,
        "gemm_ms_per_step": gemm_ms, "share_of_step": gemm_ms / ms_step if world == 1 else None,
        "kernels_ms_per_step": {k: round(v, 4) for k, v in sorted(per_tag.items())},
        "gemm_only_ms_per_step": gemm_only_ms,
        "note": "achieved = algorithmic 6*T*H*V FLOP per step / summed CUDA-event time of EVERY kernel of the step "
                "(tcgen05 GEMMs + merge + finalize + token loss + dlogits pass + tail reduce)",
    }


def vocab_parallel_record(args, dev, rank, world, tm: Timer, peaks) -> dict:
    """N>1: BASELINE config 4 (Qwen2.5-32B head, H=5120, V=152064) with the vocab rows sharded over the N ranks -- strong
    scaling of ONE micro-batch -- for both transports, each gated on parity against the unsharded fp32 reference, next to
    the unsharded step timed on one GPU of the same box in the same run (rank 0)."""
    import torch
    K, W = min(args.steps, 24), 3
    rec = {"workload": CONFIGS["c4"][7], "n_gpus": world, "steps": K, "scaling": "strong", "bwd_mode": args.mode}
    T = CONFIGS["c4"][2]
    # unsharded on one GPU (rank 0 alone computes; the others wait at the barrier)
    one_ms = 0.0
    if rank == 0:
        wl1 = DeviceWorkload("c4", dev, 0, 1, mode=args.mode)
        for i in range(W):
            wl1.step(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(8):
            wl1.step(i)
        e1.record()
        torch.cuda.synchronize()
        one_ms = e0.elapsed_time(e1) / 8
        del wl1
        torch.cuda.empty_cache()
    tm.barrier()
    one_ms = tm.max_over_ranks(one_ms)
    rec["unsharded_1gpu_ms_per_step"] = one_ms
    for ex in ("symm", "nccl"):
        wl = DeviceWorkload("c4", dev, rank, world, vocab_parallel=True, exchange=ex, exchange_ctas=args.exchange_ctas, mode=args.mode,
                            keep_full=not args.no_parity)
        name = "nvswitch_multicast" if wl.exchange is not None else "nccl"
        if ex == "symm" and wl.exchange is None:
            rec["nvswitch_multicast"] = {"unavailable": "no NVSwitch multicast mapping for this group"}
            del wl
            continue
        r = {}
        if not args.no_parity:
            r["parity"] = wl.parity()
            wl.weight_full = None
            torch.cuda.empty_cache()
        ms_total, per_tag, n_launch, metrics, _, _ = tm.time_steps(wl, K, W)
        ms = ms_total / K
        kern = sum(per_tag.values())
        r.update(ms_per_step=ms, tokens_per_s=T / (ms * 1e-3), speedup_vs_unsharded_1gpu=one_ms / ms if ms > 0 else None,
                 efficiency=one_ms / ms / world if ms > 0 else None, own_kernels_ms_per_step=kern, outside_own_kernels_ms=ms - kern,
                 kernels_ms_per_step={k: round(v, 4) for k, v in sorted(per_tag.items())}, gpu_launches_per_step=n_launch / K,
                 frac_of_sustained_peak_per_gpu=6.0 * (T - 1) * wl.H * wl.Vloc / (ms * 1e-3) / 1e12 / peaks["sustained"])
        rec[name] = r
        del wl
        torch.cuda.empty_cache()
        tm.barrier()
    return rec


def other_configs_record(args, dev, tm, peaks) -> dict:
    """N=1: the other BASELINE configs that fit one GPU (C1 0.5B shape, C2 1.5B shape, C4 32B shape unsharded), each gated on
    the same parity check and timed for a short run, so that one bench line carries every single-GPU configuration."""
    import torch
    out = {}
    for name in ("c1", "c2", "c4"):
        if name == args.config:
            continue
        try:
            wl = DeviceWorkload(name, dev, 0, 1, mode=args.mode, keep_full=True)
            par = None if args.no_parity else wl.parity()
            K = min(args.steps, 16)
            ms_total, per_tag, n_launch, metrics, _, _ = tm.time_steps(wl, K, 3)
            r = roofline_record(per_tag, ms_total / K, wl.T, wl.H, wl.Vloc, peaks, 1)
            out[name] = {"workload": wl.desc, "value": wl.T * K / (ms_total * 1e-3), "unit": "tokens/s", "ms_per_step": ms_total / K, "steps": K,
                         "frac": r["frac"], "frac_of_burst_peak": r["frac_of_burst_peak"], "achieved_tflops": r["achieved"],
                         "kernels_ms_per_step": r["kernels_ms_per_step"],
                         "parity": None if par is None else {k: par[k] for k in ("lp_rel", "ent_rel", "dH_rel", "dW_rel", "ok")}}
            del wl
        except Exception as e:  # noqa: BLE001  -- a side record must never take the bench line down
            out[name] = {"unavailable": repr(e)[:200]}
        torch.cuda.empty_cache()
    return out


def weight_broadcast_record(dev, rank, world) -> dict:
    """N>1: BASELINE config 5 -- Qwen2.5-7B bf16 parameters (15.2 GB, 339 tensors) from trainer rank 0 to the N-1 other GPUs
    (the receive_weight_update path, vllm1.py:81-94): the reference's per-parameter loop, the bucketed NCCL path and the
    scatter + relay kernels, every variant byte-checked on every receiver (tools/bench_weight_update.py has the full sweep)."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from bench_weight_update import run_sweep
    try:
        lines = run_sweep(dev, rank, world, buckets_mb=(64, 256), multicast_mb=(), scatter_mb=(256,), iters=2, scatter_p2p_mb=())
    except Exception as e:  # noqa: BLE001  -- a side record must never take the bench line down
        return {"unavailable": repr(e)[:300]}
    rec = {"workload": "C5: Qwen2.5-7B bf16 parameters, 339 tensors, trainer rank 0 -> %d receivers" % (world - 1), "unit": "GB/s",
           "nvlink_egress_peak_GBps": 900.0, "variants": {}}
    for ln in lines:
        rec["variants"][ln["variant"]] = {k: ln[k] for k in ("value", "ms", "byte_exact", "n_broadcasts", "n_transfers", "unavailable") if k in ln}
    ok = [v for v in rec["variants"].values() if v.get("byte_exact")]
    if ok:
        best = max(ok, key=lambda v: v["value"])
        rec["best"] = {"variant": [k for k, v in rec["variants"].items() if v is best][0], "value": best["value"],
                       "frac_of_nvlink_egress": best["value"] / 900.0}
        ref = rec["variants"].get("reference_per_parameter")
        if ref and ref.get("value"):
            rec["best"]["speedup_vs_reference_loop"] = best["value"] / ref["value"]
    return rec


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
        if os.environ.get("NCCL_DEBUG", "").upper() in ("WARN", "VERSION"):
            del os.environ["NCCL_DEBUG"]   # those levels print NCCL's version banner on stdout: rank 0's stdout is the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    for kv in args.tunable:
        name, val = kv.split("=")
        _abi.set_tunable(name, int(val))

    tm = Timer(dev, world)
    barrier, max_over_ranks = tm.barrier, tm.max_over_ranks
    wl = DeviceWorkload(args.config, dev, rank, world, vocab_parallel=args.parallelism == "vocab", exchange=args.exchange,
                        exchange_ctas=args.exchange_ctas, mode=args.mode, keep_full=not args.no_parity)
    H, V, T = wl.H, wl.V, wl.T
    K, W = args.steps, max(args.warmup, 3)
    peaks = load_peaks()

    # ---------------- parity gate at the benchmarked shape, before anything is timed ----------------
    parity = None
    if not args.no_parity:
        parity = wl.parity()
        if wl.vocab_parallel:
            wl.weight_full = None
        torch.cuda.empty_cache()
        if not parity["ok"]:
            raise SystemExit(f"bench.py: parity gate FAILED at the benchmarked shape: {json.dumps(parity)}")

    # ---------------- device-resident value ----------------
    with ClockSampler(local) as clocks:  # started before the warm-up: nvidia-smi needs a moment to come up
        ms_total, per_tag, n_launch, metrics, t_begin, t_end = tm.time_steps(wl, K, W)
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
    roofline = roofline_record(per_tag, ms_step, T, H, wl.Vloc, peaks, world)
    traffic, traffic_note = None, "no ncu capture for this workload"
    for rnd in ("r02", "r01"):   # DRAM bytes of the step's kernels from the committed ncu capture (C3, saved mode, unsharded only)
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", rnd, "traffic.json")))
            if tj["workload"] == f"{args.config}/{args.mode}/{args.parallelism}":
                traffic = float(sum(tj["dram_bytes_per_launch"].values()))
                traffic_note = ("dram__bytes_read+write summed over the step's four large kernels (one launch each per step), " + tj["source"]
                                + "; algorithmic operand+output bytes of the same kernels: %.1f GB" % (sum(tj["algorithmic_bytes_per_launch"].values()) / 1e9))
                break
        except Exception:  # noqa: BLE001
            pass
    roofline["traffic"], roofline["traffic_note"] = traffic, traffic_note
    launches_per_step = n_launch / K if K else 0

    # ---------------- the other backward mode on the same workload (`recompute` = the north-star design) ----------------
    other = None
    if not args.no_recompute and not wl.vocab_parallel:
        other_mode = "recompute" if args.mode == "saved" else "saved"
        wl.set_mode(other_mode)
        torch.cuda.empty_cache()
        K2 = min(K, 32)
        ms2, per_tag2, n_launch2, _, _, _ = tm.time_steps(wl, K2, W)
        r2 = roofline_record(per_tag2, ms2 / K2, T, H, wl.Vloc, peaks, world)
        other = {"bwd_mode": other_mode, "value": T * world * K2 / (ms2 * 1e-3), "unit": "tokens/s", "ms_per_step": ms2 / K2, "steps": K2,
                 "frac": r2["frac"], "frac_of_burst_peak": r2["frac_of_burst_peak"], "achieved_tflops": r2["achieved"],
                 "executed_tflops": r2["executed_tflops"], "kernels_ms_per_step": r2["kernels_ms_per_step"],
                 "scratch_bytes": ("0 B kept between forward and backward; bf16 dlogits scratch inside the backward: %d" if other_mode == "recompute"
                                   else "bf16 logits kept between forward and backward: %d") % (-(-wl.Vloc // 256) * 256 * 2 * (T - 1)),
                 "gpu_launches_per_step": n_launch2 / K2}
        wl.set_mode(args.mode)

    # ---------------- end to end through the host-buffer C entry point ----------------
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2e = None
    if not args.no_e2e and not wl.vocab_parallel:
        _abi.set_tunable("save_logits", 0 if args.mode == "recompute" else 1)
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
                      "two streams: step i+1 is submitted before the host waits for step i's metrics, so one step's dW wave tail "
                      "overlaps the next forward and e2e can come out a shade above the single-stream `value`"}
        del hs
        torch.cuda.empty_cache()

    # ---------------- the reference's torch ops on the SAME GPU (rank 0, N=1 only; SURVEY 8d "like-for-like") ----------------
    torch_gpu, api = None, None
    if rank == 0 and world == 1 and not args.no_torch_baseline:
        api = time_rl_step_api(wl, dev, min(K, 32))
        torch_gpu = time_torch_eager_gpu(wl, dev)
        if torch_gpu and "value" in torch_gpu:
            torch_gpu["speedup_of_this_repo"] = value / torch_gpu["value"]
        same_box = time_cublas_same_box(dev)
        if "bf16_tflops_sustained" in same_box:
            same_box["frac_of_it"] = roofline["achieved"] / same_box["bf16_tflops_sustained"]
        roofline["same_box_cublas"] = same_box

    # ---------------- CPU baseline (rank 0, N=1 only): the reference's own rl_step on the host cores ----------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        arm = CpuArm(args.config, wl.hidden[0].cpu(), wl.weight.cpu(), wl.cpu_batches[0])
        r = arm.run(T, 2, 1, args.cpu_budget_s or 30.0, args.cpu_sample_tokens)
        cpu = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "tokens_per_step_sample", "same_config")}
        del arm

    # ---------------- N=1: the advantage step (populate_rl_data) against the reference's pandas path on the host cores ----------------
    adv = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            from bench_advantage import run as advantage_run
            adv = advantage_run(groups=64, attempts=16, tokens=2048, iters=2)     # 1024 rollouts, ~1.6 M tokens
        except Exception as e:  # noqa: BLE001
            adv = {"unavailable": repr(e)[:200]}

    # ---------------- N=1: the other single-GPU configurations, short runs ----------------
    others = None
    if world == 1 and not args.no_other_configs:
        wl.saved = None
        torch.cuda.empty_cache()
        others = other_configs_record(args, dev, tm, peaks)

    # ---------------- N>1: config 4 vocab-parallel (the path that communicates) ----------------
    vp = None
    if world > 1 and not wl.vocab_parallel and not args.no_vocab_record:
        config_desc = wl.desc
        del wl
        torch.cuda.empty_cache()
        vp = vocab_parallel_record(args, dev, rank, world, tm, peaks)
        wl = None
    wb = None
    if world > 1 and not args.no_weight_record and args.parallelism == "dp":
        if wl is not None:
            config_desc = wl.desc
            del wl
            wl = None
        torch.cuda.empty_cache()
        wb = weight_broadcast_record(dev, rank, world)

    if rank == 0:
        is_vp = wl.vocab_parallel if wl is not None else False
        desc = wl.desc if wl is not None else config_desc
        ex = (wl.exchange is not None) if wl is not None else False
        wbytes = V * H * 2
        line = {
            "metric": "grpo_logprob_loss_fwd_bwd_tokens_per_sec", "value": value, "unit": "tokens/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if is_vp else "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": desc, "H": H, "V": V, "T": T, "parallelism": ("vocab" if is_vp else "dp") + str(world),
                       **({"exchange": "nvswitch multicast kernels (SymmetricExchange)" if ex else "nccl"} if is_vp else {}),
                       "grad_weight": "f32 accumulated across steps", "l2": "inputs exceed L2 (lm_head weight %.2f GB)" % (wbytes / 1e9),
                       "bwd_mode": args.mode,
                       "bwd_mode_note": ("saved: the forward keeps bf16 exp(logit/T - row reference) until the backward: %.2f GB at this T, so 'logits never "
                                         "materialised' holds for the `recompute` record, not for this headline; with entropy_bonus == 0 (this config) "
                                         "the backward's two GEMMs read those panels directly -- no dlogits pass" % ((-(-V // 256) * 512 * (T - 1)) / 1e9))
                       if args.mode == "saved" else "recompute: nothing of size T x V survives the forward",
                       "tunables": args.tunable, "tokens_per_gpu_per_sec": value / world if not is_vp else value},
            "parity": parity, "roofline": roofline, ("recompute" if args.mode == "saved" else "saved"): other,
            "cpu_baseline": cpu, "torch_gpu_baseline": torch_gpu, "rl_step_api": api, "e2e": e2e,
            "advantage_step": adv, "other_configs": others, "vocab_parallel": vp, "weight_broadcast": wb, "gpu_launches": int(round(launches_per_step * K)),
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
