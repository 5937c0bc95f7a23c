#!/usr/bin/env python
"""Advantage step (populate_rl_data, pipelinerl/finetune/rl/__init__.py:380-501): this repo (host scalars + K4 on the GPU)
against the reference's own pandas implementation (oracle/_ref) on the box's host cores, same list-of-dict dataset.

    python tools/bench_advantage.py [--groups 256] [--attempts 16] [--tokens 2048] [--iters 3]

Prints one JSON line: rollouts/s and tokens/s for both, and checks the outputs are equal as Python floats (f64, bit-exact).
"""
from __future__ import annotations

import argparse
import copy
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def make_dataset(groups, attempts, tokens, seed=0):
    rng = np.random.default_rng(seed)
    rows = []
    for g in range(groups):
        for a in range(attempts):
            n = int(rng.integers(tokens // 2, tokens + 1))
            k = int(rng.integers(8, max(9, n // 4)))
            ids = rng.integers(3, 150000, n).tolist()
            if rng.random() < 0.9:
                ids[-1] = 2  # eos
            r = float(rng.integers(0, 2))
            rows.append({"group_id": f"g{g}", "rollout_index": a, "step_index": 0, "input_ids": ids,
                         "labels": [-100] * k + ids[k:], "rewards": [r] * n, "advantages": [0.0] * n, "overflow": [0] * n,
                         "group_tokens": [0] * n, "num_labels": [0] * n, "old_logprobs": [0.0] * n, "ref_logprobs": [0.0] * n})
    return rows


def run(groups=256, attempts=16, tokens=2048, iters=3) -> dict:
    """Time both implementations on the same dataset; outputs must be EQUAL as Python floats (f64 bit-exact)."""
    from pipelinerl_swe_b200 import RLConfig, populate_rl_data
    data = make_dataset(groups, attempts, tokens)
    n_rows, n_tok = len(data), sum(len(r["input_ids"]) for r in data)
    cfg = RLConfig(divide_advantage_by_std=True)
    try:   # the reference's own pandas path (oracle/_ref: the reference's files, staged by oracle/build_ref.py)
        from oracle import build_ref
        rl = build_ref.load()
        ref_fn, ref_cfg, kind = rl.populate_rl_data, rl.RLConfig(divide_advantage_by_std=True), "reference (pandas groupby / merge / apply)"
    except Exception as e:  # noqa: BLE001
        import oracle
        ref_fn, ref_cfg, kind = oracle.oracle_populate_rl_data, oracle.OracleRLConfig(divide_advantage_by_std=True), f"port ({e!r})"[:120]

    def timed(fn):
        best, out = 1e30, None
        for _ in range(iters):
            d = copy.deepcopy(data)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = fn(d)
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        return best, out

    populate_rl_data(copy.deepcopy(data[:32]), 2, cfg)  # warm the library
    t_ours, ours = timed(lambda d: populate_rl_data(d, 2, cfg))
    t_ref, ref = timed(lambda d: ref_fn(d, 2, ref_cfg))
    same = all(a[k] == b[k] for a, b in zip(ours, ref) for k in ("advantages", "group_tokens", "overflow", "num_labels"))
    return {"metric": "populate_rl_data", "rollouts": n_rows, "tokens": n_tok, "ours_s": round(t_ours, 4),
            "reference_s": round(t_ref, 4), "reference_kind": kind, "speedup": round(t_ref / t_ours, 2),
            "ours_tokens_per_s": round(n_tok / t_ours), "reference_tokens_per_s": round(n_tok / t_ref),
            "host_cores": os.cpu_count(), "outputs_bit_identical": bool(same),
            "note": "both include the list-of-dict in/out conversion the reference's contract requires; ours = host scalars + K4 on the GPU + D2H"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--groups", type=int, default=256)
    ap.add_argument("--attempts", type=int, default=16)
    ap.add_argument("--tokens", type=int, default=2048)
    ap.add_argument("--iters", type=int, default=3)
    args = ap.parse_args()
    print(json.dumps(run(args.groups, args.attempts, args.tokens, args.iters)))


if __name__ == "__main__":
    main()
