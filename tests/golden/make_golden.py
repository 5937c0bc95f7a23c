"""Generate golden vectors by running the UNMODIFIED reference (authoring container only).

Usage:  python tests/golden/make_golden.py          (needs /root/reference)

Imports ``pipelinerl.finetune.rl`` etc. straight from /root/reference (nothing is
copied), runs them on small seeded inputs and stores inputs + outputs under
tests/golden/ as .npz / .json.  The GPU box has no /root/reference, so tests only read
the committed fixtures.  Two import stubs are needed for modules the container lacks
(omegaconf, accelerate); they provide names only, no behaviour used by the path.
"""
from __future__ import annotations

import json
import os
import sys
import types
from importlib.machinery import ModuleSpec

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("GRPO_REFERENCE", "/root/reference")


def import_reference():
    import datasets  # noqa: F401  (must precede the stubs)
    import transformers  # noqa: F401

    for name, attrs in (("omegaconf", {"DictConfig": dict}), ("accelerate", {"Accelerator": type("Accelerator", (), {})})):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                m = types.ModuleType(name)
                m.__spec__ = ModuleSpec(name, None)
                for k, v in attrs.items():
                    setattr(m, k, v)
                sys.modules[name] = m
    sys.path.insert(0, REF)
    import pipelinerl.finetune.rl as rl
    import pipelinerl.finetune.rl.utils as rlu
    from pipelinerl.finetune.types import PipelineBatchEncoding
    import pipelinerl.finetune.data as data
    import pipelinerl.finetune.utils as futils

    return rl, rlu, PipelineBatchEncoding, data, futils


class StubModel(torch.nn.Module):
    """A 'model' that is only its lm_head: logits = hidden @ W.T (fp32 math on bf16 values)."""

    def __init__(self, hidden, weight, values=None):
        super().__init__()
        self.hidden = torch.nn.Parameter(hidden.float())
        self.weight = torch.nn.Parameter(weight.float())
        if values is not None:
            self.value_head = True
            self.values = torch.nn.Parameter(values.float())

    def forward(self, input_ids, attention_mask, labels, position_ids=None, **kw):
        h = self.hidden if self.hidden.dim() == 3 else self.hidden[None]
        out = types.SimpleNamespace(logits=h @ self.weight.t())
        if hasattr(self, "values"):
            out.value = self.values
        return out


def make_batch(g, lens, V, n_prompt, packed=True, start_pos=0, rows=1):
    """Synthetic packed row (or [rows, L] unpacked batch) with all RL fields."""
    T = sum(lens)
    ids = torch.randint(0, V, (rows, T), generator=g)
    labels = ids.clone()
    pos = torch.zeros(rows, T, dtype=torch.long)
    rewards = torch.zeros(rows, T)
    adv = torch.zeros(rows, T)
    gtok = torch.zeros(rows, T)
    nlab = torch.zeros(rows, T)
    ovf = torch.zeros(rows, T)
    for r in range(rows):
        o = 0
        for i, n in enumerate(lens):
            pos[r, o:o + n] = torch.arange(n) + (start_pos if i == 0 else 0)
            k = min(n_prompt, n - 1)
            labels[r, o:o + k] = -100
            if i > 0:
                labels[r, o] = -100
            rewards[r, o:o + n] = float(torch.randint(0, 4, (1,), generator=g)) * 0.5 - 0.5
            adv[r, o:o + n] = torch.randn(1, generator=g).item()
            gtok[r, o:o + n] = float(sum(lens)) / len(lens)
            nlab[r, o:o + n] = float((labels[r, o:o + n] != -100).sum())
            ovf[r, o:o + n] = float(i % 3 == 2)
            o += n
    m = (labels != -100).float()
    old = -torch.rand(rows, T, generator=g) * 3 * m
    ref = -torch.rand(rows, T, generator=g) * 3 * m
    b = dict(input_ids=ids, attention_mask=torch.ones_like(ids), labels=labels, rewards=rewards, advantages=adv,
             ref_logprobs=ref, old_logprobs=old, group_tokens=gtok, num_labels=nlab, overflow=ovf,
             model_version=0, is_packed=packed)
    if packed:
        b["position_ids"] = pos
        b["seq_boundaries"] = torch.tensor(np.concatenate([[0], np.cumsum(lens)]), dtype=torch.int32)
    return b


RL_CASES = {
    # name: (lens, H, V, n_prompt, packed, start_pos, rows, cfg kwargs, (step, max_step), value_head)
    "ppo_kl_packed": ([17, 20, 11], 32, 257, 5, True, 0, 1, dict(policy_loss="ppo", kl_coef=0.1, final_kl_coef=0.1, batch_size=4, temperature=0.7, clamp_log_ratio_ref_new_value=5, epsilon=0.2), (3, 10), False),
    "reinforce_gn_of_ent": ([9, 30, 14, 7], 48, 300, 3, True, 0, 1, dict(policy_loss="reinforce", kl_coef=0.05, final_kl_coef=0.2, entropy_bonus=0.01, final_entropy_bonus=0.03, group_normalization=True, overlong_filtering=True, batch_size=8, epsilon=4, temperature=1.0), (7, 20), False),
    "ppo_rewards_relu": ([15, 12, 9, 12], 16, 130, 4, True, 0, 1, dict(policy_loss="ppo", use_advantages=False, relu_log_p_weights=True, kl_coef=0.0, final_kl_coef=0.0, batch_size=2, epsilon=0.1, clamp_log_ratio_ref_new_value=0.5), (0, 5), False),
    "sp_slice_midseq": ([13, 19], 32, 200, 2, True, 40, 1, dict(policy_loss="ppo", batch_size=4, kl_coef=0.1, final_kl_coef=0.1), (1, 4), False),
    "unpacked_rows": ([16], 24, 150, 6, False, 0, 3, dict(policy_loss="ppo", batch_size=3, kl_coef=0.1, final_kl_coef=0.1), (2, 9), False),
    "value_head": ([12, 18], 16, 120, 3, True, 0, 1, dict(policy_loss="ppo", batch_size=2, kl_coef=0.1, final_kl_coef=0.1, value_loss_coef=0.5), (1, 3), True),
}


def tens_to_np(b):
    return {k: (v.numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in b.items()}


def gen_rl_step(rl, PBE):
    for name, (lens, H, V, n_prompt, packed, start_pos, rows, cfgkw, (step, max_step), vh) in RL_CASES.items():
        g = torch.Generator().manual_seed(sum(map(ord, name)))
        b = make_batch(g, lens, V, n_prompt, packed, start_pos, rows)
        T = sum(lens)
        hidden = torch.randn(rows, T, H, generator=g).bfloat16()
        weight = (torch.randn(V, H, generator=g) * 0.3).bfloat16()
        values = torch.rand(rows, T, generator=g) if vh else None
        if name != "ppo_rewards_relu":
            # old/ref logprobs near the current policy so un-clipped branches are exercised too
            z = (hidden.float() @ weight.float().t()) / cfgkw.get("temperature", 1.0)
            lp = torch.log_softmax(z, -1)[:, :-1].gather(2, b["input_ids"][:, 1:, None]).squeeze(2)
            m = (b["labels"][:, 1:] != -100).float()
            b["old_logprobs"][:, 1:] = (lp + 0.15 * torch.randn(lp.shape, generator=g)) * m
            b["ref_logprobs"][:, 1:] = (lp + 0.5 * torch.randn(lp.shape, generator=g)) * m
        model = StubModel(hidden if rows > 1 or not packed else hidden[0], weight, values)
        cfg = rl.RLConfig(**cfgkw)
        loss, stats = rl.rl_step(model, PBE(**b), step, max_step, cfg)
        loss.backward()
        out = tens_to_np(b)
        out.update(hidden=hidden.float().numpy(), weight=weight.float().numpy(),
                   loss=loss.detach().numpy(), grad_hidden=model.hidden.grad.numpy().reshape(rows, T, H),
                   grad_weight=model.weight.grad.numpy())
        if vh:
            out.update(values=values.numpy(), grad_values=model.values.grad.numpy())
        np.savez_compressed(os.path.join(HERE, f"rl_step_{name}.npz"), **out)
        meta = dict(cfg=cfgkw, step=step, max_step=max_step, stats=stats, value_head=vh)
        with open(os.path.join(HERE, f"rl_step_{name}.json"), "w") as f:
            json.dump(meta, f, indent=1, sort_keys=True)
        print(name, float(loss), len(stats))


def gen_sentinel(rl, PBE, futils):
    sb = futils.create_sentinel_batch("cpu", tokenizer=None, model_version=7)
    g = torch.Generator().manual_seed(5)
    hidden = torch.randn(8, 16, generator=g).bfloat16()
    weight = torch.randn(50, 16, generator=g).bfloat16()
    model = StubModel(hidden, weight)
    loss, stats = rl.rl_step(model, sb, 0, 10, rl.RLConfig(batch_size=4))
    fields = {k: getattr(sb, k) for k in type(sb).model_fields if getattr(sb, k) is not None}
    out = tens_to_np(fields)
    out.update(hidden=hidden.float().numpy(), weight=weight.float().numpy(), loss=loss.detach().numpy())
    np.savez_compressed(os.path.join(HERE, "rl_step_sentinel.npz"), **out)
    with open(os.path.join(HERE, "rl_step_sentinel.json"), "w") as f:
        json.dump(dict(cfg=dict(batch_size=4), step=0, max_step=10, stats=stats, value_head=False), f, indent=1)
    print("sentinel", float(loss), stats)


def rollout_rows(g, n_groups, attempts, eos, rng_reward, steps=1, singleton=False):
    rows = []
    for gi in range(n_groups):
        n_att = 1 if (singleton and gi == 0) else attempts
        for a in range(n_att):
            for s in range(steps):
                n = int(torch.randint(4, 12, (1,), generator=g))
                ids = torch.randint(3, 50, (n,), generator=g).tolist()
                if (gi + a) % 3 != 0:
                    ids[-1] = eos
                n_prompt = int(torch.randint(1, 3, (1,), generator=g))
                labels = [-100] * n_prompt + ids[n_prompt:]
                reward = rng_reward(gi, a)
                rows.append(dict(group_id=f"g{gi}", rollout_index=a, step_index=s, input_ids=ids, labels=labels,
                                 rewards=[reward] * n, advantages=[0.0] * n, overflow=[0] * n, group_tokens=[0] * n,
                                 num_labels=[1 if t != -100 else 0 for t in labels],
                                 old_logprobs=[0.0] * n, ref_logprobs=[0.0] * n, model_version=gi))
    return rows


def gen_advantages(rl, data):
    import copy
    cases = {}
    g = torch.Generator().manual_seed(11)
    cases["binary_4x4_std"] = (rollout_rows(g, 4, 4, 2, lambda gi, a: float((gi * 3 + a * 5) % 3 == 0)), dict(divide_advantage_by_std=True))
    cases["real_3x5_nostd"] = (rollout_rows(g, 3, 5, 2, lambda gi, a: 0.37 * gi - 0.11 * a * a), dict(divide_advantage_by_std=False))
    cases["singleton_and_steps"] = (rollout_rows(g, 3, 3, 2, lambda gi, a: float(a) * 0.5 + gi, steps=2, singleton=True), dict(divide_advantage_by_std=True))
    out = {}
    for name, (rows, cfgkw) in cases.items():
        inp = copy.deepcopy(rows)
        res = rl.populate_rl_data(rows, 2, rl.RLConfig(**cfgkw))
        out[name] = dict(cfg=cfgkw, eos=2, input=inp,
                         output=[{k: r[k] for k in ("advantages", "group_tokens", "overflow", "num_labels")} for r in res])
        print(name, len(rows))
    # collate_packed on the first case's output (seq_parallel 4 forces a sentinel pad)
    tok = type("Tok", (), {"eos_token_id": 2})()
    ex = [{k: r[k] for k in ("input_ids", "labels", "rewards", "advantages", "overflow", "group_tokens", "num_labels", "old_logprobs", "ref_logprobs", "model_version")} for r in cases["binary_4x4_std"][0][:5]]
    pb = data.collate_packed(copy.deepcopy(ex), tok, 4)
    fields = {k: getattr(pb, k) for k in type(pb).model_fields if getattr(pb, k) is not None}
    out["collate_packed"] = dict(examples=ex, seq_parallel=4, eos=2,
                                 batch={k: (v.tolist() if isinstance(v, torch.Tensor) else v) for k, v in fields.items()})
    with open(os.path.join(HERE, "advantages.json"), "w") as f:
        json.dump(out, f)


def gen_aggregate(rlu):
    stats = {"loss": [0.5, -0.25, 1.0], "max_kl": [0.1, 0.7, 0.3], "min_reward": [0.0, 1.0, 0.5], "ratio_new_old_sum": [10.0, 12.5],
             "entropy": [2.0, 4.0, 6.0], "max_loss": [0.5, 1.0], "num_output_tokens_sum": [100, 200]}
    res = rlu.aggregate_rl_stats(stats, 5)
    with open(os.path.join(HERE, "aggregate.json"), "w") as f:
        json.dump(dict(input=stats, num_samples=5, output=res), f, indent=1)


def gen_final_norm():
    """Final RMSNorm (transformers Qwen2RMSNorm, unmodified) + the reference's ValueHead (value_model.py:40-52), bf16, with
    autograd gradients for an upstream (dy, dvalue)."""
    from transformers.models.qwen2.modeling_qwen2 import Qwen2RMSNorm
    from pipelinerl.finetune import value_model as vm  # the reference's module, after import_reference()'s stubs
    out = {}
    for name, (T, H) in {"a": (37, 64), "b": (5, 264)}.items():
        g = torch.Generator().manual_seed(100 + T)
        norm = Qwen2RMSNorm(H, eps=1e-6)
        with torch.no_grad():
            norm.weight.copy_(1.0 + 0.2 * torch.randn(H, generator=g))
        head = vm.ValueHead(H)
        with torch.no_grad():
            head.output.weight.copy_(0.05 * torch.randn(1, H, generator=g))
            head.output.bias.fill_(0.125)
        norm, head = norm.to(torch.bfloat16), head.to(torch.bfloat16)
        x = (torch.randn(T, H, generator=g) * 3.0).to(torch.bfloat16).requires_grad_(True)
        y = norm(x)
        value = head(y)
        dy = (torch.randn(T, H, generator=g) * 0.1).to(torch.bfloat16)
        dv = (torch.randn(T, generator=g) * 0.1).to(torch.bfloat16)
        torch.autograd.backward([y, value], [dy, dv])
        f = lambda t: t.detach().float().numpy()  # noqa: E731  (bf16 values are exact in f32)
        out.update({f"{name}_x": f(x), f"{name}_gamma": f(norm.weight), f"{name}_vw": f(head.output.weight), f"{name}_vb": f(head.output.bias),
                    f"{name}_y": f(y), f"{name}_value": f(value), f"{name}_dy": f(dy), f"{name}_dv": f(dv), f"{name}_dx": f(x.grad),
                    f"{name}_dgamma": f(norm.weight.grad), f"{name}_dvw": f(head.output.weight.grad), f"{name}_dvb": f(head.output.bias.grad)})
    np.savez_compressed(os.path.join(HERE, "final_norm.npz"), **out)


if __name__ == "__main__":
    torch.set_num_threads(1)  # deterministic reduction order
    if sys.argv[1:] == ["final_norm"]:  # only the final-norm fixture (leaves the other files untouched)
        import_reference()
        gen_final_norm()
        sys.exit(0)
    rl, rlu, PBE, data, futils = import_reference()
    gen_final_norm()
    gen_rl_step(rl, PBE)
    gen_sentinel(rl, PBE, futils)
    gen_advantages(rl, data)
    gen_aggregate(rlu)
