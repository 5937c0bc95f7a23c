"""Randomised parity of the whole drop-in ``rl_step`` (+ ``loss.backward()``) against the REFERENCE's own ``rl_step``.

TEST INFRASTRUCTURE (only tests/ run it).  The comparator is ``pipelinerl.finetune.rl.rl_step`` from ``oracle/_ref`` -- the
reference's unmodified files (rl/__init__.py:130-377) staged by ``oracle/build_ref.py`` -- on a model that is only its lm_head
in fp32 (the same bf16 values, fp32 math: SURVEY 8c "primary" numerics); when ``oracle/_ref`` is not staged the oracle
restatement takes its place and the record says so.

    python tests/fuzz_rl_step.py <first_seed> <n_seeds> [cases_per_seed]     (GPU box; prints one line per seed)

Each case draws: packed row or unpacked rows, sequence lengths, prompt lengths (some sequences fully masked), a slice that starts
mid-sequence, H, V (ragged / tiny / several panels), temperature, every RLConfig switch (ppo / reinforce, advantages or rewards,
relu weights, epsilon, KL coefficient + clamp, entropy bonus, group normalisation, overlong filtering, value head + value loss),
the opt-in switches (skip masked tokens, fused f32 gradient accumulation + fold),
step / max_step for the decayed coefficients, hot (repeated) targets, confident policies (the target is the peak of the row, p up to
0.999+: gradients are differences of nearly equal numbers), and the saved / recompute backward.  Tolerances are the
north star's: loss 1e-3, every statistic 2e-3 (+2e-4 absolute), gradients 2e-2 of the largest reference entry.
"""
from __future__ import annotations

import math
import os
import random
import sys
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

LOSS_RTOL, STAT_RTOL, STAT_ATOL, GRAD_RTOL = 1e-3, 2e-3, 2e-4, 2e-2


def _reference():
    """(kind, rl module or oracle module, PipelineBatchEncoding or None)"""
    try:
        from oracle import build_ref
        rl = build_ref.load()
        return "reference", rl, sys.modules["pipelinerl.finetune.types"].PipelineBatchEncoding
    except Exception:  # noqa: BLE001
        import oracle
        return "port", oracle, None


class _RefLM(torch.nn.Module):
    """logits = hidden @ W^T in fp32 on the bf16 values (+ the value-head output the reference reads at :239-246)."""

    def __init__(self, hidden, weight, values):
        super().__init__()
        self.hidden = torch.nn.Parameter(hidden.float())
        self.weight = torch.nn.Parameter(weight.float())
        if values is not None:
            self.value_head = True
            self.values = torch.nn.Parameter(values.float())

    def forward(self, **kw):
        out = types.SimpleNamespace(logits=torch.nn.functional.linear(self.hidden, self.weight))
        if hasattr(self, "values"):
            out.value = self.values
        return out


class _OurLM(torch.nn.Module):
    def __init__(self, hidden, weight, values):
        super().__init__()
        self.hidden = torch.nn.Parameter(hidden.clone())
        V, H = weight.shape
        self.lm_head = torch.nn.Linear(H, V, bias=False, dtype=torch.bfloat16)
        with torch.no_grad():
            self.lm_head.weight.copy_(weight)
        if values is not None:
            self.value_head = True
            self.values = torch.nn.Parameter(values.float().clone())

    def forward(self, **kw):
        out = types.SimpleNamespace(logits=self.lm_head(self.hidden))
        if hasattr(self, "values"):
            out.value = self.values
        return out


def draw_case(rnd: random.Random, seed: int):
    from helpers import near_policy_logprobs, synthetic_packed_batch
    packed = rnd.random() < 0.75
    H = rnd.choice([16, 64, 136, 520])
    V = rnd.choice([40, 130, 257, 1000, 1032, 5000])
    if packed:
        lens = [rnd.choice([2, 3, 9, 17, 40, 129, 300]) for _ in range(rnd.randint(1, 7))]
        rows = 1
    else:
        lens = [rnd.choice([6, 21, 64])]
        rows = rnd.randint(2, 4)
    n_prompt = rnd.choice([0, 1, 4, 1000])          # 1000: everything but the last token of each sequence is masked
    start_pos = rnd.choice([0, 0, 5]) if packed else 0   # a sequence-parallel slice that starts mid-sequence (:159-163)
    batch = synthetic_packed_batch(lens, V, seed=seed, n_prompt=n_prompt, rows=rows, packed=packed, start_pos=start_pos)
    T = sum(lens)
    if rnd.random() < 0.3:                          # hot tokens: many positions share a target
        pool = torch.randint(0, V, (3,), generator=torch.Generator().manual_seed(seed))
        ids = pool[torch.randint(0, 3, batch["input_ids"].shape, generator=torch.Generator().manual_seed(seed + 1))]
        batch["labels"] = torch.where(batch["labels"] == -100, batch["labels"], ids)
        batch["input_ids"] = ids
    if rnd.random() < 0.2 and T > 4:                # a hole of masked tokens inside a completion
        batch["labels"][:, T // 2] = -100
    g = torch.Generator().manual_seed(seed + 7)
    hidden = torch.randn(rows, T, H, generator=g).bfloat16()
    weight = (torch.randn(V, H, generator=g) * rnd.choice([0.02, 0.1, 0.3])).bfloat16()
    temp = rnd.choice([0.6, 1.0, 1.3])
    if rnd.random() < 0.25:                         # a confident policy: the next token is the peak of most rows (p_target up to 0.999+)
        wn = weight.float() / (weight.float().norm(dim=1, keepdim=True) + 1e-6)
        weight = wn.bfloat16()
        nxt = torch.roll(batch["input_ids"], shifts=-1, dims=1)
        # margin ln(V) + 2..6 nats: 1 - p_target between ~0.13 and ~0.0025.  (Closer to 1 the comparison itself is ill-conditioned:
        # two fp32 evaluations of 1 - p differ by 1e-5 / (1 - p) through the summation order of the logits alone.)
        margin = math.log(V) + rnd.uniform(2.0, 6.0)
        hidden = (wn[nxt] * margin * temp + 0.3 * torch.randn(rows, T, H, generator=g)).bfloat16()
    if rnd.random() < 0.7:
        near_policy_logprobs(batch, hidden, weight, temp, seed=seed, s_old=rnd.choice([0.02, 0.3]), s_ref=rnd.choice([0.05, 0.5, 3.0]))
    value_head = rnd.random() < 0.2
    ent = rnd.choice([0.0, 0.0, 0.01])
    kl = rnd.choice([0.0, 0.1, 0.02])
    cfg = dict(policy_loss=rnd.choice(["ppo", "ppo", "reinforce"]), use_advantages=rnd.random() < 0.8, epsilon=rnd.choice([0.1, 0.2, 4.0]),
               batch_size=rnd.choice([1, 4, 4096]), kl_coef=kl, final_kl_coef=rnd.choice([kl, 0.3]), entropy_bonus=ent,
               final_entropy_bonus=rnd.choice([ent, 0.02]), relu_log_p_weights=rnd.random() < 0.2,
               clamp_log_ratio_ref_new_value=rnd.choice([0.3, 5.0, 10.0]), overlong_filtering=rnd.random() < 0.3,
               group_normalization=rnd.random() < 0.4, temperature=temp, value_loss_coef=rnd.choice([0.0, 0.5]) if value_head else 0.0)
    max_step = rnd.choice([1, 10, 100])
    step = rnd.randint(0, max_step)
    values = torch.rand(rows, T, generator=g) if value_head else None
    return dict(batch=batch, hidden=hidden, weight=weight, values=values, cfg=cfg, step=step, max_step=max_step,
                recompute=rnd.random() < 0.3, skip_masked=rnd.random() < 0.25, fused_acc=rnd.random() < 0.25, desc=dict(packed=packed, lens=lens, rows=rows, H=H, V=V, n_prompt=n_prompt, start_pos=start_pos,
                                                       value_head=value_head, step=(step, max_step), **cfg))


def run_reference(case, dev):
    kind, mod, PBE = _reference()
    b = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in case["batch"].items()}
    model = _RefLM(case["hidden"].to(dev), case["weight"].to(dev), None if case["values"] is None else case["values"].to(dev))
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        if kind == "reference":
            pbe = PBE(**case["batch"], model_version=0).to_device(dev)   # validated on the host, then moved (finetune_loop.py:109)
            loss, stats = mod.rl_step(model, pbe, case["step"], case["max_step"], mod.RLConfig(**case["cfg"]))
        else:
            loss, stats = mod.oracle_rl_step(model.hidden[0] if b["is_packed"] else model.hidden, model.weight, b, case["step"], case["max_step"],
                                             mod.OracleRLConfig(**case["cfg"]), values=getattr(model, "values", None))
        if loss.requires_grad:
            loss.backward()
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    g = lambda p: None if p.grad is None else p.grad.detach()  # noqa: E731
    return kind, float(loss.detach()), stats, g(model.hidden), g(model.weight), (g(model.values) if hasattr(model, "values") else None)


def run_ours(case, dev):
    from pipelinerl_swe_b200 import (RLConfig, fold_main_grad, rl_step, set_fused_grad_accumulation, set_save_logits,
                                     set_skip_masked_tokens)
    b = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in case["batch"].items()}
    model = _OurLM(case["hidden"], case["weight"], case["values"]).to(dev)
    set_save_logits(not case["recompute"])
    set_skip_masked_tokens(bool(case.get("skip_masked")))          # opt-in switches: same loss, statistics and gradients
    set_fused_grad_accumulation(bool(case.get("fused_acc")))
    try:
        loss, stats = rl_step(model, b, case["step"], case["max_step"], RLConfig(**case["cfg"]))
        if loss.requires_grad:
            loss.backward()
        if case.get("fused_acc"):
            fold_main_grad(model)                                   # f32 main_grad -> .grad, what an unmodified optimizer reads
    finally:
        set_save_logits(True)
        set_skip_masked_tokens(False)
        set_fused_grad_accumulation(False)
    g = lambda p: None if p.grad is None else p.grad.detach()  # noqa: E731
    return float(loss.detach()), stats, g(model.hidden), g(model.lm_head.weight), (g(model.values) if hasattr(model, "values") else None)


def _rel(a, b):
    return float((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-20))


def compare(case, dev) -> tuple[bool, str, str]:
    kind, rloss, rstats, rgh, rgw, rgv = run_reference(case, dev)
    loss, stats, gh, gw, gv = run_ours(case, dev)
    why = []
    if abs(loss - rloss) > LOSS_RTOL * abs(rloss) + 1e-5:
        why.append(f"loss {loss} vs {rloss}")
    if set(stats) != set(rstats):
        why.append(f"stat keys differ: {sorted(set(stats) ^ set(rstats))}")
    else:
        for k, w in rstats.items():
            if not abs(stats[k] - w) <= STAT_ATOL + STAT_RTOL * abs(w):
                why.append(f"stat {k}: {stats[k]} vs {w}")
    for name, a, r in (("dH", gh, rgh), ("dW", gw, rgw), ("dvalues", gv, rgv)):
        if r is None or float(r.abs().max()) == 0.0:
            if a is not None and float(a.float().abs().max()) != 0.0:
                why.append(f"{name}: reference gradient is zero / absent, ours is not")
            continue
        if a is None:
            why.append(f"{name}: missing")
        elif _rel(a.reshape(r.shape), r) > GRAD_RTOL:
            why.append(f"{name} rel {_rel(a.reshape(r.shape), r):.3e}")
    return not why, "; ".join(why[:4]), kind


def fuzz(seed: int, n_cases: int = 12, dev="cuda:0"):
    rnd = random.Random(seed)
    bad, kind = [], "?"
    for c in range(n_cases):
        case = draw_case(rnd, seed * 1000 + c)
        try:
            ok, why, kind = compare(case, dev)
        except AssertionError as e:   # the reference asserts (non-finite, group_tokens <= 0 ...): ours must raise as well, not this harness
            ok, why = False, f"assertion in one arm only: {e!r}"[:200]
        if not ok:
            bad.append((c, why, case["desc"]))
    return not bad, f"seed {seed}: {n_cases} cases vs {kind}, {len(bad)} bad {bad[:2]}"


if __name__ == "__main__":
    first, n = int(sys.argv[1]), int(sys.argv[2])
    cases = int(sys.argv[3]) if len(sys.argv) > 3 else 12
    n_bad = 0
    for s in range(first, first + n):
        ok, msg = fuzz(s, cases)
        n_bad += not ok
        print("FUZZ", "OK" if ok else "FAIL", msg, flush=True)
    print(f"FUZZ SUMMARY seeds {first}..{first + n - 1} x {cases} cases: {n_bad} failing seeds")
    sys.exit(1 if n_bad else 0)


# ---- error behaviour: the reference raises (rl/__init__.py:209,222,237,264,283,313); the drop-in must raise the same type ----------
def error_cases():
    """(name, mutate(case) -> None) pairs; each makes the reference's rl_step raise."""
    def gt_zero(c):
        c["cfg"]["group_normalization"] = True
        c["batch"]["group_tokens"][:, 3] = 0.0
    def ref_inf(c):
        c["batch"]["ref_logprobs"][:, -1] = float("inf")
    def ref_nan_on_masked_token(c):      # the reference's isfinite().all() covers masked positions too (:237)
        c["batch"]["labels"][:, 1] = -100
        c["batch"]["ref_logprobs"][:, 1] = float("nan")
    def hidden_nan(c):
        c["hidden"][:, 2, 0] = float("nan")
    def unknown_policy(c):
        c["cfg"]["policy_loss"] = "grpo_v9"
    def loss_overflow(c):
        c["cfg"].update(policy_loss="ppo", use_advantages=True, relu_log_p_weights=False)
        c["batch"]["advantages"][:, -1] = float("inf")
        c["batch"]["labels"][:, -1] = c["batch"]["input_ids"][:, -1]
    return [("group_tokens_zero", gt_zero), ("ref_logprob_inf", ref_inf), ("ref_logprob_nan_masked", ref_nan_on_masked_token),
            ("hidden_nan", hidden_nan), ("unknown_policy_loss", unknown_policy), ("non_finite_loss", loss_overflow)]


def check_error_case(name, mutate, dev="cuda:0"):
    """Returns (reference exception type, our exception type)."""
    out = []
    for run in (run_reference, run_ours):
        case = draw_case(random.Random(77), 77000)
        case["cfg"].update(group_normalization=False, overlong_filtering=False)
        case["values"] = None
        case["cfg"]["value_loss_coef"] = 0.0
        mutate(case)
        try:
            run(case, dev)
            out.append(None)
        except Exception as e:  # noqa: BLE001
            out.append(type(e))
    return tuple(out)


# ---- autograd usage patterns a trainer may apply to the returned loss ----------------------------------------------------------
def usage_patterns(dev="cuda:0"):
    """Each pattern is run on the reference arm and on the drop-in with identical inputs; returns [(name, ok, detail)]."""
    kind, mod, PBE = _reference()
    from pipelinerl_swe_b200 import RLConfig, rl_step
    out = []

    def both(case):
        ref_model = _RefLM(case["hidden"].to(dev), case["weight"].to(dev), None)
        our_model = _OurLM(case["hidden"], case["weight"], None).to(dev)
        b = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in case["batch"].items()}

        def ref_step():
            if kind == "reference":
                return mod.rl_step(ref_model, PBE(**case["batch"], model_version=0).to_device(dev), case["step"], case["max_step"], mod.RLConfig(**case["cfg"]))
            return mod.oracle_rl_step(ref_model.hidden[0] if b["is_packed"] else ref_model.hidden, ref_model.weight, b, case["step"], case["max_step"],
                                      mod.OracleRLConfig(**case["cfg"]))

        def our_step():
            return rl_step(our_model, b, case["step"], case["max_step"], RLConfig(**case["cfg"]))
        return ref_model, our_model, ref_step, our_step

    def grads_close(ref_model, our_model):
        e1 = _rel(our_model.lm_head.weight.grad, ref_model.weight.grad)
        e2 = _rel(our_model.hidden.grad.reshape(ref_model.hidden.grad.shape), ref_model.hidden.grad)
        return max(e1, e2) < GRAD_RTOL, f"dW {e1:.2e} dH {e2:.2e}"

    def base_case(seed, **cfg):
        case = draw_case(random.Random(seed), seed)
        case["values"] = None
        case["cfg"].update(value_loss_coef=0.0, **cfg)
        return case

    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        # 1. scaled loss (gradient accumulation: DeepSpeed / accelerate divide the loss before backward)
        rm, om, rs, os_ = both(base_case(901))
        (rs()[0] * 0.25).backward(); (os_()[0] * 0.25).backward()
        out.append(("scaled_loss", *grads_close(rm, om)))
        # 2. two micro-batches' losses summed, ONE backward (both graphs alive at once; entropy term on)
        rm, om, rs, os_ = both(base_case(902, entropy_bonus=0.01, final_entropy_bonus=0.01))
        (rs()[0] + 2.0 * rs()[0]).backward(); (os_()[0] + 2.0 * os_()[0]).backward()
        out.append(("two_graphs_one_backward", *grads_close(rm, om)))
        # 3. the same graph backpropagated twice (retain_graph): the saved exponentials are consumed by the first pass
        rm, om, rs, os_ = both(base_case(903))
        lr, lo = rs()[0], os_()[0]
        lr.backward(retain_graph=True); lr.backward()
        lo.backward(retain_graph=True); lo.backward()
        out.append(("backward_twice", *grads_close(rm, om)))
        # 4. evaluation: no_grad -> same loss and stats, nothing to backpropagate
        rm, om, rs, os_ = both(base_case(904))
        with torch.no_grad():
            (lr, sr), (lo, so) = rs(), os_()
        ok = (not lo.requires_grad) and abs(float(lo) - float(lr)) <= LOSS_RTOL * abs(float(lr)) + 1e-5 and set(so) == set(sr)
        out.append(("no_grad", ok, f"loss {float(lo)} vs {float(lr)}"))
        # 5. torch.autograd.grad with respect to the hidden states only (no weight gradient wanted)
        rm, om, rs, os_ = both(base_case(905))
        (gr,) = torch.autograd.grad(rs()[0], rm.hidden)
        (go,) = torch.autograd.grad(os_()[0], om.hidden)
        e = _rel(go.reshape(gr.shape), gr)
        out.append(("grad_wrt_hidden_only", e < GRAD_RTOL and om.lm_head.weight.grad is None, f"dH {e:.2e}"))
        # 6. a side stream
        rm, om, rs, os_ = both(base_case(906))
        rs()[0].backward()
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            os_()[0].backward()
        torch.cuda.current_stream(dev).wait_stream(side)
        out.append(("side_stream", *grads_close(rm, om)))
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    return out
