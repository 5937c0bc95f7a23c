"""Shared test helpers: golden loading and synthetic batches (no reference imports)."""
from __future__ import annotations

import glob
import json
import os

import numpy as np
import torch

from oracle import OracleRLConfig

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TENSOR_KEYS = ("input_ids", "attention_mask", "labels", "position_ids", "rewards", "advantages", "ref_logprobs",
               "old_logprobs", "group_tokens", "num_labels", "overflow", "seq_boundaries")


def golden_rl_cases() -> list[str]:
    return sorted(os.path.basename(p)[len("rl_step_"):-len(".npz")] for p in glob.glob(os.path.join(GOLDEN, "rl_step_*.npz")))


def load_rl_case(name: str):
    z = np.load(os.path.join(GOLDEN, f"rl_step_{name}.npz"))
    meta = json.load(open(os.path.join(GOLDEN, f"rl_step_{name}.json")))
    batch = {k: torch.from_numpy(z[k]) for k in TENSOR_KEYS if k in z.files}
    batch["is_packed"] = bool(z["is_packed"])
    cfg = OracleRLConfig(**meta["cfg"])
    extra = {k: torch.from_numpy(z[k]) for k in ("hidden", "weight", "loss", "grad_hidden", "grad_weight", "values", "grad_values") if k in z.files}
    return batch, cfg, meta, extra


def assert_stats_close(got: dict, want: dict, rtol=1e-3, atol=1e-4):
    assert set(got.keys()) == set(want.keys()), (sorted(set(got) ^ set(want)))
    for k, w in want.items():
        g = got[k]
        assert abs(g - w) <= atol + rtol * abs(w), f"stat {k}: got {g} want {w}"
