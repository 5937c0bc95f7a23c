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


# ---- synthetic inputs shared by GPU tests, smoke() and bench.py --------------------------------
def synthetic_packed_batch(lens, V, seed=0, n_prompt=8, rows=1, packed=True, start_pos=0):
    """A packed [1,T] row (or [rows, L] unpacked batch) with every RL field, seeded, CPU tensors."""
    g = torch.Generator().manual_seed(seed)
    T = int(sum(lens))
    ids = torch.randint(0, V, (rows, T), generator=g)
    labels = ids.clone()
    pos = torch.zeros(rows, T, dtype=torch.long)
    f = lambda: torch.zeros(rows, T)  # noqa: E731
    rewards, adv, gtok, nlab, ovf = f(), f(), f(), f(), f()
    for r in range(rows):
        o = 0
        for i, n in enumerate(lens):
            pos[r, o:o + n] = torch.arange(n) + (start_pos if i == 0 else 0)
            k = min(n_prompt, n - 1)
            labels[r, o:o + k] = -100
            rewards[r, o:o + n] = float(torch.randint(0, 2, (1,), generator=g))
            adv[r, o:o + n] = torch.randn(1, generator=g).item()
            gtok[r, o:o + n] = float(T) / len(lens)
            nlab[r, o:o + n] = float(max(n - k, 1))
            ovf[r, o:o + n] = float(i % 5 == 4)
            o += n
    m = (labels != -100).float()
    old = -(torch.rand(rows, T, generator=g) * 2 + 0.1) * m
    ref = -(torch.rand(rows, T, generator=g) * 2 + 0.1) * m
    b = dict(input_ids=ids, attention_mask=torch.ones_like(ids), labels=labels, rewards=rewards, advantages=adv,
             ref_logprobs=ref, old_logprobs=old, group_tokens=gtok, num_labels=nlab, overflow=ovf, is_packed=packed)
    if packed:
        b["position_ids"] = pos
    return b


def near_policy_logprobs(batch, hidden, weight, temperature, seed=0, s_old=0.05, s_ref=0.1):
    """old/ref logprobs = current policy logprob + noise on labelled positions (SURVEY 8d recipe)."""
    g = torch.Generator().manual_seed(seed)
    h = hidden.float().reshape(batch["input_ids"].shape[0], -1, hidden.shape[-1])
    lp = []
    for r in range(h.shape[0]):
        z = (h[r] @ weight.float().t()) / temperature
        lp.append(torch.log_softmax(z, -1)[:-1].gather(1, batch["input_ids"][r, 1:, None]).squeeze(1))
    lp = torch.stack(lp)
    m = (batch["labels"][:, 1:] != -100).float()
    batch["old_logprobs"][:, 1:] = (lp + s_old * torch.randn(lp.shape, generator=g)) * m
    batch["ref_logprobs"][:, 1:] = (lp + s_ref * torch.randn(lp.shape, generator=g)) * m
    return batch


class StubLM(torch.nn.Module):
    """A 'model' reduced to what the hot path touches: fixed hidden states + an lm_head.

    forward(**inputs) -> namespace(logits=self.lm_head(hidden)[, value]) like an HF causal LM,
    so rl_step's lm_head swap is exercised exactly as with a real model.
    """

    def __init__(self, H, V, T, seed=0, rows=1, w_std=0.05, value_head=False, dtype=torch.bfloat16):
        super().__init__()
        g = torch.Generator().manual_seed(seed)
        self.hidden = torch.nn.Parameter(torch.randn(rows, T, H, generator=g).to(dtype))
        self.lm_head = torch.nn.Linear(H, V, bias=False, dtype=dtype)
        with torch.no_grad():
            self.lm_head.weight.copy_((torch.randn(V, H, generator=g) * w_std).to(dtype))
        if value_head:
            self.value_head = True
            self.values = torch.nn.Parameter(torch.rand(rows, T, generator=g))

    def forward(self, input_ids=None, attention_mask=None, position_ids=None, labels=None, **kw):
        import types
        out = types.SimpleNamespace(logits=self.lm_head(self.hidden))
        if hasattr(self, "values"):
            out.value = self.values
        return out


class StubTokenizer:
    eos_token_id = 2
    padding_side = "right"


def synthetic_rollout_examples(lens, V, seed=0, n_prompt=3, per_token_rewards=False):
    """Examples as populate_rl_data leaves them (per-token lists; per-rollout columns constant inside a rollout)."""
    g = torch.Generator().manual_seed(seed)
    out = []
    for i, n in enumerate(lens):
        ids = torch.randint(3, V, (n,), generator=g).tolist()
        k = min(n_prompt, n)
        labels = [-100] * k + ids[k:]
        lp = lambda: [0.0] * k + (-(torch.rand(n - k, generator=g) * 3)).tolist()  # noqa: E731
        rew = float(torch.randint(0, 2, (1,), generator=g))
        rewards = [rew] * n
        if per_token_rewards and n > 1:
            rewards[-1] = rew + 0.5
        out.append({
            "input_ids": ids, "labels": labels, "attention_mask": [1] * n, "model_version": 3 + (i % 2),
            "rewards": rewards, "advantages": [float(torch.randn(1, generator=g))] * n, "old_logprobs": lp(), "ref_logprobs": lp(),
            "group_tokens": [float(37 + i)] * n, "num_labels": [float(n - k)] * n, "overflow": [float(i % 3 == 2)] * n,
        })
    return out
