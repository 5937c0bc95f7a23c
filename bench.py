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
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS))
    ap.add_argument("--parallelism", default="dp", choices=["dp", "vocab"])
    ap.add_argument("--mode", default="auto", choices=["auto", "recompute", "fused"], help="backward strategy")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-tokens", type=int, default=0, help="tokens per CPU-baseline call (0 = auto)")
    ap.add_argument("--no-e2e", action="store_true")
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
    w = weight.float().requires_grad_(True)
    loss, stats = oracle.oracle_rl_step(h, w, sub, 0, 1, cfg)
    loss.backward()
    return float(loss)


def time_cpu_reference(hidden_cpu, weight_cpu, batch, ocfg, n_tokens, steps, warmup):
    import torch
    torch.set_num_threads(os.cpu_count() or 1)
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
        self.index, self.proc, self.lines = index, None, []

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
            self.lines.append(line.strip())

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
