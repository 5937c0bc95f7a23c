"""aggregate_rl_stats: consumer of rl_step's stats keys (pipelinerl/finetune/rl/utils.py:8-22).

The reduction is chosen by SUBSTRING of the key, in the reference's order: "min" -> min,
"max" -> max, key == "loss" -> sum, "sum" -> sum, otherwise sum / num_samples.  Values are
reduced as float32, like ``torch.Tensor(v)`` in the reference.
"""
from __future__ import annotations

import numpy as np


def aggregate_rl_stats(rl_stats: dict, num_samples: int) -> dict[str, float]:
    out: dict[str, float] = {}
    for k, v in rl_stats.items():
        a = np.asarray(v, dtype=np.float32)
        if "min" in k:
            r = a.min()
        elif "max" in k:
            r = a.max()
        elif k == "loss" or "sum" in k:
            r = a.sum(dtype=np.float32)
        else:
            r = a.sum(dtype=np.float32) / np.float32(num_samples)
        out["rl/" + k] = float(r)
    return out
