"""GPU parity for the HBM-bound kernels: K3 token loss, segment bounds, K4 advantages, K6 buckets, and
the host-buffer C entry point.  Integer / index / byte work must be bit-exact."""
import copy
import ctypes as C
import json
import os

import numpy as np
import pytest
import torch

import oracle
from helpers import GOLDEN, StubLM, assert_stats_close, near_policy_logprobs, synthetic_packed_batch

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda:0")


def _to_dev(batch):
    return {k: (v.to(_dev()) if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}


# ---- segment boundaries: bit-exact (rl/__init__.py:159-170) ------------------------------------
@pytest.mark.parametrize("T,p_zero,first", [(1, 0.0, 5), (9, 0.3, 0), (1024, 0.01, 7), (1025, 0.5, 0), (8192, 0.002, 3),
                                            (1 << 20, 0.001, 0), (5000, 0.0, 9), (4097, 1.0, 0)])
def test_segment_bounds_bit_exact(T, p_zero, first):
    from pipelinerl_swe_b200 import ops
    g = torch.Generator().manual_seed(T)
    pos = torch.randint(1, 100, (T,), generator=g)
    pos[torch.rand(T, generator=g) < p_zero] = 0
    pos[0] = first
    want, n = oracle.oracle_segments(pos.clone())
    got = ops.segment_bounds(pos.to(_dev()))
    assert got.dtype == torch.int32
    assert got.cpu().tolist() == want.tolist()
    assert got.numel() == n + 1


# ---- K3: loss, 34 stats and the gradient coefficients vs the oracle -----------------------------
K3_CASES = {
    "ppo_kl": dict(policy_loss="ppo", kl_coef=0.1, final_kl_coef=0.1, batch_size=64, epsilon=0.2, clamp_log_ratio_ref_new_value=0.3),
    "ppo_gn_of_ent": dict(policy_loss="ppo", kl_coef=0.02, final_kl_coef=0.3, entropy_bonus=0.01, final_entropy_bonus=0.02,
                          group_normalization=True, overlong_filtering=True, batch_size=8, epsilon=0.05),
    "reinforce": dict(policy_loss="reinforce", kl_coef=0.1, final_kl_coef=0.1, batch_size=16, epsilon=0.1),
    "reinforce_rewards_relu": dict(policy_loss="reinforce", use_advantages=False, relu_log_p_weights=True, kl_coef=0.0,
                                   final_kl_coef=0.0, batch_size=4),
}


@pytest.mark.parametrize("name", list(K3_CASES))
@pytest.mark.parametrize("packed,lens,rows", [(True, [700, 33, 1, 2, 1500, 260], 1), (False, [257], 5), (True, [8192], 1)])
def test_token_loss_vs_oracle(name, packed, lens, rows):
    from pipelinerl_swe_b200 import RLConfig
    from pipelinerl_swe_b200 import rl as prl
    cfgkw = K3_CASES[name]
    batch = synthetic_packed_batch(lens, 1000, seed=len(name) + rows, rows=rows, packed=packed)
    L = batch["labels"].shape[1]
    g = torch.Generator().manual_seed(1)
    lp = -(torch.rand(rows, L - 1, generator=g) * 2 + 0.05)
    ent = torch.rand(rows, L - 1, generator=g) * 3
    m = (batch["labels"][:, 1:] != -100).float()
    batch["old_logprobs"][:, 1:] = (lp + 0.2 * torch.randn(lp.shape, generator=g)) * m
    batch["ref_logprobs"][:, 1:] = (lp + 0.4 * torch.randn(lp.shape, generator=g)) * m
    ocfg = oracle.OracleRLConfig(**cfgkw)
    lp_r, ent_r = lp.clone().requires_grad_(True), ent.clone().requires_grad_(True)
    want_loss, want_stats = oracle.oracle_token_loss(lp_r, ent_r, batch, 3, 10, ocfg)
    want_loss.backward()

    cfg = prl.make_loss_cfg(RLConfig(**cfgkw), 3, 10, packed, False)
    db = _to_dev(batch)
    lp_d, ent_d = lp.to(_dev()).requires_grad_(True), ent.to(_dev()).requires_grad_(True)
    sink = {}
    loss = prl.TokenLossFunction.apply(lp_d, ent_d, None, db, cfg, sink)
    loss.backward()
    metr = sink["metrics"].tolist()
    stats = prl.stats_from_metrics(metr, cfg, batch["input_ids"].numel(), False,
                                   (oracle.oracle_linear_decay_coef(3, 10, ocfg.kl_coef, ocfg.final_kl_coef),
                                    oracle.oracle_linear_decay_coef(3, 10, ocfg.entropy_bonus, ocfg.final_entropy_bonus)))
    assert abs(loss.item() - want_loss.item()) <= 1e-4 * abs(want_loss.item()) + 1e-6
    assert_stats_close(stats, want_stats, rtol=2e-4, atol=1e-5)
    # integer-valued stats are exact
    for k in ("num_output_tokens_sum", "num_nans", "input_size"):
        assert stats[k] == want_stats[k]
    torch.testing.assert_close(lp_d.grad.cpu(), lp_r.grad, rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(ent_d.grad.cpu(), ent_r.grad, rtol=1e-4, atol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("lens,poison", [([2], False), ([1, 1, 2], False), ([3, 1, 5], True), ([64, 1, 1, 1, 30], True)])
def test_token_loss_tiny_rows_and_masked_nonfinite(lens, poison):
    """Shortest possible rows (one shifted position, one-token segments) and NaN / inf rewards + advantages at MASKED positions:
    `mask_sum` multiplies by the mask and then nan_to_num's (rl/utils.py:28-30), so they must vanish from every sum while
    `num_nans` still counts them (rl/__init__.py:345), exactly as the oracle does."""
    from pipelinerl_swe_b200 import RLConfig
    from pipelinerl_swe_b200 import rl as prl
    cfgkw = dict(policy_loss="ppo", kl_coef=0.1, final_kl_coef=0.1, batch_size=4, epsilon=0.2)
    batch = synthetic_packed_batch(lens, 1000, seed=sum(lens), n_prompt=1)
    L = batch["labels"].shape[1]
    g = torch.Generator().manual_seed(3)
    lp = -(torch.rand(1, L - 1, generator=g) + 0.05)
    ent = torch.rand(1, L - 1, generator=g)
    m = (batch["labels"][:, 1:] != -100).float()
    batch["old_logprobs"][:, 1:] = (lp + 0.1 * torch.randn(lp.shape, generator=g)) * m
    batch["ref_logprobs"][:, 1:] = (lp + 0.2 * torch.randn(lp.shape, generator=g)) * m
    if poison:
        masked = (batch["labels"][0] == -100).nonzero().flatten().tolist()
        assert masked
        batch["rewards"][0, masked[0]] = float("nan")
        batch["advantages"][0, masked[0]] = float("inf")
        batch["advantages"][0, masked[-1]] = float("nan")
    ocfg = oracle.OracleRLConfig(**cfgkw)
    want_loss, want_stats = oracle.oracle_token_loss(lp.clone(), ent.clone(), batch, 0, 1, ocfg)
    cfg = prl.make_loss_cfg(RLConfig(**cfgkw), 0, 1, True, False)
    sink = {}
    loss = prl.TokenLossFunction.apply(lp.to(_dev()), ent.to(_dev()), None, _to_dev(batch), cfg, sink)
    stats = prl.stats_from_metrics(sink["metrics"].tolist(), cfg, batch["input_ids"].numel(), False, (0.1, 0.0))
    assert set(stats) == set(want_stats)
    if "loss" in want_stats:
        assert abs(loss.item() - want_loss.item()) <= 1e-5 * abs(want_loss.item()) + 1e-7
        assert_stats_close(stats, want_stats, rtol=2e-4, atol=1e-6)
        assert stats["num_nans"] == want_stats["num_nans"] and stats["num_output_tokens_sum"] == want_stats["num_output_tokens_sum"]


def test_token_loss_flags_nonfinite():
    from pipelinerl_swe_b200 import RLConfig
    from pipelinerl_swe_b200 import rl as prl
    batch = synthetic_packed_batch([40, 24], 100, seed=2)
    L = batch["labels"].shape[1]
    lp = -torch.rand(1, L - 1)
    lp[0, 30] = float("nan")
    cfg = prl.make_loss_cfg(RLConfig(batch_size=2), 0, 1, True, False)
    sink = {}
    prl.TokenLossFunction.apply(lp.to(_dev()), torch.zeros(1, L - 1, device=_dev()), None, _to_dev(batch), cfg, sink)
    with pytest.raises(AssertionError):
        prl.stats_from_metrics(sink["metrics"].tolist(), cfg, L, False)


def test_rl_step_requires_cuda_batch():
    from pipelinerl_swe_b200 import RLConfig, rl_step
    from pipelinerl_swe_b200._abi import B200GrpoError
    batch = synthetic_packed_batch([20, 12], 100, seed=2)
    model = StubLM(32, 100, 32)
    with pytest.raises(B200GrpoError):
        rl_step(model, batch, 0, 1, RLConfig(batch_size=2))


# ---- rl_step on a mid-size packed row vs the oracle (tests the whole K1->K3->K2 chain) ------------
@pytest.mark.parametrize("policy", ["ppo", "reinforce"])
def test_rl_step_midsize_vs_oracle(policy):
    from pipelinerl_swe_b200 import RLConfig, rl_step
    H, V, lens = 256, 32000, [300, 211, 129, 384]
    T = sum(lens)
    batch = synthetic_packed_batch(lens, V, seed=11, n_prompt=32)
    model = StubLM(H, V, T, seed=5, w_std=0.05)
    cfgkw = dict(policy_loss=policy, batch_size=4, kl_coef=0.1, final_kl_coef=0.05, temperature=0.9, epsilon=0.02,
                 entropy_bonus=0.001, final_entropy_bonus=0.001)
    near_policy_logprobs(batch, model.hidden.detach()[0], model.lm_head.weight.detach(), 0.9, seed=3)
    hidden = model.hidden.detach().float()[0].requires_grad_(True)
    weight = model.lm_head.weight.detach().float().requires_grad_(True)
    want_loss, want_stats = oracle.oracle_rl_step(hidden, weight, batch, 2, 10, oracle.OracleRLConfig(**cfgkw))
    want_loss.backward()
    model = model.to(_dev())
    loss, stats = rl_step(model, _to_dev(batch), 2, 10, RLConfig(**cfgkw))
    loss.backward()
    assert abs(loss.item() - want_loss.item()) <= 1e-3 * abs(want_loss.item()) + 1e-6
    assert_stats_close(stats, want_stats, rtol=2e-3, atol=2e-4)
    gw, gh = model.lm_head.weight.grad.float().cpu(), model.hidden.grad.float().cpu()[0]
    assert float((gw - weight.grad).abs().max() / weight.grad.abs().max()) < 2e-2
    assert float((gh - hidden.grad).abs().max() / hidden.grad.abs().max()) < 2e-2


# ---- K4: advantages vs the reference's pandas output (golden) and the oracle -------------------
ADV = json.load(open(os.path.join(GOLDEN, "advantages.json")))


@pytest.mark.parametrize("name", [k for k in ADV if k != "collate_packed"])
def test_populate_rl_data_vs_reference_golden(name):
    from pipelinerl_swe_b200 import RLConfig, populate_rl_data
    case = ADV[name]
    rows = copy.deepcopy(case["input"])
    res = populate_rl_data(rows, case["eos"], RLConfig(**case["cfg"]))
    assert res is rows
    for got, want in zip(res, case["output"]):
        assert got["num_labels"] == want["num_labels"]          # integer: exact
        assert got["overflow"] == want["overflow"]              # 0.0 / 1.0: exact
        assert got["group_tokens"] == want["group_tokens"]      # f64 Python floats, bit-identical to pandas' mean
        assert got["advantages"] == want["advantages"]          # f64, bit-identical: K4 follows pandas' Kahan mean / Welford variance


def test_populate_rl_data_large_vs_oracle():
    from pipelinerl_swe_b200 import RLConfig, populate_rl_data
    g = np.random.default_rng(3)
    rows = []
    for gi in range(64):
        for a in range(16):
            n = int(g.integers(8, 400))
            ids = g.integers(3, 1000, n).tolist()
            if (gi + a) % 4:
                ids[-1] = 2
            npmt = int(g.integers(1, 6))
            labels = [-100] * npmt + ids[npmt:]
            rows.append(dict(group_id=f"g{gi}", rollout_index=a, step_index=0, input_ids=ids, labels=labels,
                             rewards=[float(g.integers(0, 2)) if gi % 3 else float(g.normal())] * n, advantages=[0.0] * n, overflow=[0] * n,
                             group_tokens=[0] * n, num_labels=[1 if t != -100 else 0 for t in labels]))
    want = oracle.oracle_populate_rl_data(copy.deepcopy(rows), 2, oracle.OracleRLConfig(divide_advantage_by_std=True))
    got = populate_rl_data(rows, 2, RLConfig(divide_advantage_by_std=True))
    for a, b in zip(got, want):
        assert a["num_labels"] == b["num_labels"] and a["overflow"] == b["overflow"]
        assert a["group_tokens"] == b["group_tokens"] and a["advantages"] == b["advantages"]   # bit-exact f64


# ---- K6: bucket pack / unpack round trip, byte-exact --------------------------------------------
@pytest.mark.parametrize("seed", [400, 401])
def test_populate_rl_data_random_datasets_bit_exact_vs_the_reference_pandas(seed):
    """Shuffled rows, single-member / zero-variance groups, odd rewards, several step_index per group: every output list == the
    reference's own populate_rl_data (oracle/_ref, pandas) as Python floats."""
    from fuzz_advantage import fuzz
    ok, msg = fuzz(seed)
    print(msg)
    assert ok, msg


def test_bucket_roundtrip_bit_exact():
    from pipelinerl_swe_b200 import ops
    d = _dev()
    g = torch.Generator(device=d).manual_seed(0)
    shapes = [(3584,), (512, 3584), (1,), (7,), (18944, 128), (0,), (129, 3), (1000003,)]
    src = [torch.randn(*s, device=d, generator=g).bfloat16() for s in shapes]
    src[4] = torch.randn(*shapes[4], device=d, generator=g)  # an fp32 master weight: cast fused into the pack
    n = ops.bucket_numel(src)
    assert n == sum((t.numel() + 7) // 8 * 8 for t in src)
    bucket = torch.zeros(n, dtype=torch.bfloat16, device=d)
    ops.bucket_pack(src, bucket)
    off = 0
    for t in src:  # wire content == param.data.bfloat16() (finetune_loop.py:247)
        want = t.bfloat16().reshape(-1)
        assert torch.equal(bucket[off:off + t.numel()].view(torch.int16), want.view(torch.int16))
        off += (t.numel() + 7) // 8 * 8
    dst = [torch.empty_like(t) for t in src]
    ops.bucket_unpack(bucket, dst)
    for a, b in zip(dst, src):
        assert torch.equal(a.bfloat16().view(torch.int16), b.bfloat16().view(torch.int16))


# ---- host-buffer C entry point == the piecewise calls -------------------------------------------
@pytest.mark.parametrize("entropy_bonus", [0.0, 0.01])
def test_rl_step_host_entry_matches_piecewise(entropy_bonus):
    """entropy_bonus != 0: both run the recompute backward (an entropy gradient is never formed from the saved exponentials); the
    host entry's dlogits scratch then spans its backward workspace and the unused saved-exponentials region."""
    from pipelinerl_swe_b200 import RLConfig, rl_step
    from pipelinerl_swe_b200 import rl as prl
    from pipelinerl_swe_b200 import host_step
    H, V, lens = 128, 5000, [100, 60, 95]
    T = sum(lens)
    batch = synthetic_packed_batch(lens, V, seed=21)
    model = StubLM(H, V, T, seed=8).to(_dev())
    cfgkw = dict(policy_loss="ppo", batch_size=3, kl_coef=0.1, final_kl_coef=0.1, temperature=1.0, entropy_bonus=entropy_bonus,
                 final_entropy_bonus=entropy_bonus)
    loss, stats = rl_step(model, _to_dev(batch), 0, 1, RLConfig(**cfgkw))
    loss.backward()
    step = host_step.HostRLStep(T, H, V, model.lm_head.weight.detach(), RLConfig(**cfgkw), 0, 1)
    step.load_batch(batch, model.hidden.detach()[0].cpu())
    m = step.run(want_grad_hidden=True)
    torch.cuda.synchronize()
    stats2 = prl.stats_from_metrics(m.tolist(), step.cfg, T, False)
    assert_stats_close(stats2, stats, rtol=1e-6, atol=1e-7)
    assert torch.equal(step.grad_weight, model.lm_head.weight.grad)
    assert torch.equal(step.grad_hidden_host.view(torch.int16)[:-1], model.hidden.grad[0].cpu().view(torch.int16)[:-1])


# ---------------------------------------------------------------------------------------------
# step metrics accumulator: device accumulation == list appends + aggregate_rl_stats
# ---------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("value_head", [False, True])
def test_step_metrics_match_reference_aggregation(value_head):
    from collections import defaultdict
    from helpers import StubLM, near_policy_logprobs, synthetic_packed_batch
    from pipelinerl_swe_b200 import RLConfig, StepMetrics, aggregate_rl_stats, rl_step, rl_step_async
    d = torch.device("cuda:0")
    H, V = 128, 2048 + 8
    cfg = RLConfig(policy_loss="ppo", batch_size=4, kl_coef=0.05, final_kl_coef=0.02, entropy_bonus=0.01, final_entropy_bonus=0.01,
                   temperature=0.9)
    sm = StepMetrics(d, has_values=value_head)
    lists, losses = defaultdict(list), []
    micro = [([40, 25, 60], 3), ([30, 30], 4), ([17, 90, 11, 9], 5)]
    for lens, seed in micro:
        T = sum(lens)
        model = StubLM(H, V, T, seed=seed, value_head=value_head).to(d)
        batch = synthetic_packed_batch(lens, V, seed=seed)
        batch = near_policy_logprobs(batch, model.hidden.detach().cpu()[0], model.lm_head.weight.detach().cpu(), cfg.temperature, seed)
        batch["seq_boundaries"] = torch.tensor([0] + list(np.cumsum(lens)))
        bd = {k: (v.to(d) if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
        loss, stats = rl_step(model, bd, 2, 10, cfg)
        for k, v in stats.items():
            lists[k].append(v)
        loss2 = rl_step_async(model, bd, 2, 10, cfg, sm)
        losses.append((loss.item(), loss2.item()))
    # a sentinel batch advances the counters only; an unlabeled micro-batch contributes input_size only
    model = StubLM(H, V, 50, seed=9, value_head=value_head).to(d)
    batch = synthetic_packed_batch([50], V, seed=9)
    batch["seq_boundaries"] = torch.tensor([0, 50])
    bd = {k: (v.to(d) if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
    rl_step_async(model, bd, 2, 10, cfg, sm, record=False)
    bd["labels"] = torch.full_like(bd["labels"], -100)
    _, stats = rl_step(model, bd, 2, 10, cfg)
    assert list(stats) == ["input_size"]
    lists["input_size"].append(stats["input_size"])
    rl_step_async(model, bd, 2, 10, cfg, sm)
    want = aggregate_rl_stats(lists, num_samples=9)
    got, counters = sm.reduce(9)
    assert all(a == b for a, b in losses)
    assert counters == {"labelled_micro_batches": 3, "micro_batches": 5, "samples": 3 + 2 + 4 + 1 + 1}
    assert set(got) == set(want)
    for k, w in want.items():
        assert abs(got[k] - w) <= 1e-5 * max(1.0, abs(w)), (k, got[k], w)
    # reduce() reset the accumulator
    got2, counters2 = sm.reduce(9)
    assert got2 == {} and counters2["micro_batches"] == 0


# ---------------------------------------------------------------------------------------------
# packed-batch ingest: wire form + device expansion == collate_packed + to_device, bit for bit
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("lens,sp", [([5, 1, 9, 0, 4], 1), ([5, 1, 9, 0, 4], 8), ([1], 1), ([700, 33, 1200, 2, 513], 4),
                                      ([3] * 4000, 1)])
def test_packed_ingest_bit_exact(lens, sp):
    from helpers import StubTokenizer, synthetic_rollout_examples
    from pipelinerl_swe_b200 import collate_packed
    from pipelinerl_swe_b200.ingest import ingest_packed, pack_rollouts
    ex = synthetic_rollout_examples(lens, 152064, seed=len(lens) + sp)
    want = collate_packed(copy.deepcopy(ex), StubTokenizer(), sp).to_device(_dev())
    wire = pack_rollouts(copy.deepcopy(ex), StubTokenizer(), sp)
    got = wire.to_device(_dev())
    torch.cuda.synchronize()
    for k in ("input_ids", "labels", "attention_mask", "position_ids", "rewards", "advantages", "ref_logprobs", "old_logprobs",
              "group_tokens", "num_labels", "overflow", "seq_boundaries"):
        a, b = getattr(got, k), getattr(want, k)
        assert a.dtype == b.dtype and a.shape == b.shape, k
        assert torch.equal(a, b), k
    assert (got.is_packed, got.padding, got.model_version) == (want.is_packed, want.padding, want.model_version)
    T = want.input_ids.shape[1]
    assert wire.nbytes < 0.3 * 60 * T + 30 * len(lens) + 200       # vs the reference's 60 B per token
    # per-token rewards cannot take the wire form: same result through the fallback
    ex2 = synthetic_rollout_examples(lens, 152064, seed=1, per_token_rewards=True)
    want2 = collate_packed(copy.deepcopy(ex2), StubTokenizer(), sp).to_device(_dev())
    got2 = ingest_packed(copy.deepcopy(ex2), StubTokenizer(), sp, _dev())
    assert torch.equal(got2.rewards, want2.rewards) and torch.equal(got2.labels, want2.labels)


@pytest.mark.parametrize("seed", [0, 1])
def test_packed_ingest_random_layouts_bit_exact(seed):
    """Random rollout lists (empty, single-token and long rollouts, any count, seq_parallel 1 / 2 / 4 / 8 padding): wire form +
    device expansion == collate_packed + to_device, every field, bit for bit."""
    import random
    from helpers import StubTokenizer, synthetic_rollout_examples
    from pipelinerl_swe_b200 import collate_packed
    from pipelinerl_swe_b200.ingest import pack_rollouts
    rnd = random.Random(seed)
    for case in range(20):
        lens = [rnd.choice([0, 1, 1, 2, 3, 7, 64, 129, 700, 2049]) for _ in range(rnd.randint(1, 40))]
        if sum(lens) == 0:
            lens.append(5)
        sp = rnd.choice([1, 2, 4, 8])
        ex = synthetic_rollout_examples(lens, 152064, seed=seed * 100 + case, n_prompt=rnd.choice([0, 3, 10000]))
        want = collate_packed(copy.deepcopy(ex), StubTokenizer(), sp).to_device(_dev())
        got = pack_rollouts(copy.deepcopy(ex), StubTokenizer(), sp).to_device(_dev())
        for k in ("input_ids", "labels", "attention_mask", "position_ids", "rewards", "advantages", "ref_logprobs", "old_logprobs",
                  "group_tokens", "num_labels", "overflow", "seq_boundaries"):
            a, b = getattr(got, k), getattr(want, k)
            assert a.dtype == b.dtype and a.shape == b.shape and torch.equal(a, b), (k, lens, sp)
        assert (got.is_packed, got.padding, got.model_version) == (want.is_packed, want.padding, want.model_version)


def test_rl_step_on_ingested_batch_equals_collated_batch():
    from helpers import StubTokenizer, synthetic_rollout_examples
    from pipelinerl_swe_b200 import RLConfig, collate_packed, rl_step
    from pipelinerl_swe_b200.ingest import ingest_packed
    lens, H, V = [120, 64, 200, 77], 128, 4096
    ex = synthetic_rollout_examples(lens, V, seed=8, n_prompt=10)
    model = StubLM(H, V, sum(lens), seed=2).to(_dev())
    cfg = RLConfig(batch_size=4, kl_coef=0.1, final_kl_coef=0.1)
    l1, s1 = rl_step(model, collate_packed(copy.deepcopy(ex), StubTokenizer(), 1).to_device(_dev()), 0, 1, cfg)
    l2, s2 = rl_step(model, ingest_packed(copy.deepcopy(ex), StubTokenizer(), 1, _dev()), 0, 1, cfg)
    assert l1.item() == l2.item() and s1 == s2


# ---------------------------------------------------------------------------------------------
# opt-in: leave positions whose next token is not labelled out of the lm_head GEMMs
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n_prompt", [40, 400])
def test_skip_masked_tokens_gives_the_same_step(n_prompt):
    from pipelinerl_swe_b200 import RLConfig, rl_step, set_skip_masked_tokens
    from pipelinerl_swe_b200 import _abi
    H, V, lens = 128, 4096 + 8, [300, 150, 420, 90]
    T = sum(lens)
    batch = synthetic_packed_batch(lens, V, seed=13, n_prompt=n_prompt)     # n_prompt=400: two rollouts are (almost) all prompt
    res = []
    for skip in (False, True):
        model = StubLM(H, V, T, seed=4).to(_dev())
        near_policy_logprobs(batch, model.hidden.detach().cpu()[0], model.lm_head.weight.detach().cpu(), 1.0, seed=2)
        cfg = RLConfig(batch_size=4, kl_coef=0.1, final_kl_coef=0.1, entropy_bonus=0.01, final_entropy_bonus=0.01)
        n0 = _abi.launch_count()
        try:
            set_skip_masked_tokens(skip)
            loss, stats = rl_step(model, _to_dev(batch), 0, 1, cfg)
            loss.backward()
        finally:
            set_skip_masked_tokens(False)
        res.append((loss.item(), stats, model.lm_head.weight.grad.float().clone(), model.hidden.grad.float().clone()))
    (l0, s0, gw0, gh0), (l1, s1, gw1, gh1) = res
    assert abs(l0 - l1) <= 1e-6 * abs(l0) + 1e-9
    assert_stats_close(s1, s0, rtol=1e-5, atol=1e-7)
    assert float((gw0 - gw1).abs().max()) <= 2e-2 * float(gw0.abs().max())      # bf16 outputs, different token-sum order
    assert float((gh0 - gh1).abs().max()) <= 2e-2 * float(gh0.abs().max())
    labelled_next = (batch["labels"][0, 1:] != -100)
    assert float(gh1[0, :-1][~labelled_next].abs().max()) == 0.0                  # skipped rows get exactly zero gradient


# ---------------------------------------------------------------------------------------------------------------
# Final RMSNorm + value head in one sweep
# ---------------------------------------------------------------------------------------------------------------
def _final_norm_ref(x, gamma, eps, vw, vb, dy, dv):
    """The oracle on the GPU tensors' CPU copies, with autograd gradients (bf16 semantics of the transformers module)."""
    import oracle
    xc, gc = x.cpu().requires_grad_(True), gamma.cpu().requires_grad_(True)
    vwc = vw.cpu().requires_grad_(True) if vw is not None else None
    vbc = vb.cpu().requires_grad_(True) if vb is not None else None
    y, value = oracle.oracle_final_norm_value(xc, gc, eps, vwc, vbc)
    outs, grads = [y], [dy.cpu()]
    if value is not None:
        outs.append(value); grads.append(dv.cpu().to(value.dtype))
    torch.autograd.backward(outs, grads)
    return y.detach(), (value.detach() if value is not None else None), xc.grad, gc.grad, (vwc.grad if vwc is not None else None), \
        (vbc.grad if vbc is not None else None)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["a", "b"])
def test_final_norm_value_golden(case):
    """CUDA kernel vs outputs of transformers' Qwen2RMSNorm + the reference's ValueHead themselves (tests/golden/final_norm.npz)."""
    import numpy as np
    from helpers import GOLDEN
    from pipelinerl_swe_b200 import ops
    d = torch.device("cuda:0")
    z = np.load(os.path.join(GOLDEN, "final_norm.npz"))
    t = lambda k: torch.from_numpy(z[f"{case}_{k}"]).to(torch.bfloat16).to(d)  # noqa: E731
    x, gamma, vw, vb = t("x"), t("gamma"), t("vw").reshape(-1), t("vb")
    y, rstd, value = ops.final_norm_fwd(x, gamma, 1e-6, vw, vb)
    # forward: every rounding point of the transformers module is reproduced; the f32 mean may differ in the last bit, which can
    # flip a bf16 rounding of the normalised value (1 bf16 ulp = 2^-8 relative)
    torch.testing.assert_close(y.float(), t("y").float(), rtol=2 ** -7, atol=1e-6)
    assert (y != t("y")).float().mean().item() < 0.01
    torch.testing.assert_close(value, t("value").float(), rtol=1e-2, atol=2e-3)   # golden value is bf16-rounded
    dx, dgamma, dvw, dvb = ops.final_norm_bwd(x, gamma, rstd, t("dy"), t("dv").float(), vw)
    for got, want, name in ((dx.float(), t("dx").float(), "dx"), (dgamma, t("dgamma").float(), "dgamma"), (dvw, t("dvw").float().reshape(-1), "dvw"),
                            (dvb, t("dvb").float(), "dvb")):
        err = (got - want).abs().max().item() / max(want.abs().max().item(), 1e-12)
        assert err < 2e-2, (name, err)   # gradients: 2e-2 of max (north-star tolerance; the golden side is bf16 autograd)


@pytest.mark.gpu
@pytest.mark.parametrize("T,H,with_value", [(1, 8, True), (300, 1536, True), (257, 3584, False), (64, 5120, True), (33, 8192, True),
                                            (1000, 896, False)])
def test_final_norm_value_vs_oracle(T, H, with_value):
    from pipelinerl_swe_b200 import ops
    d = torch.device("cuda:0")
    g = torch.Generator(device=d).manual_seed(T * 7 + H)
    x = (torch.randn(T, H, device=d, generator=g) * 2.5).bfloat16()
    gamma = (1.0 + 0.3 * torch.randn(H, device=d, generator=g)).bfloat16()
    vw = (0.05 * torch.randn(H, device=d, generator=g)).bfloat16() if with_value else None
    vb = torch.full((1,), 0.25, device=d).bfloat16() if with_value else None
    dy = (0.1 * torch.randn(T, H, device=d, generator=g)).bfloat16()
    dv = 0.1 * torch.randn(T, device=d, generator=g) if with_value else None
    y, rstd, value = ops.final_norm_fwd(x, gamma, 1e-6, vw, vb)
    dx, dgamma, dvw, dvb = ops.final_norm_bwd(x, gamma, rstd, dy, dv, vw)
    ry, rvalue, rdx, rdg, rdvw, rdvb = _final_norm_ref(x, gamma, 1e-6, vw, vb, dy, dv.bfloat16() if dv is not None else None)
    torch.testing.assert_close(y.float().cpu(), ry.float(), rtol=2 ** -7, atol=1e-6)
    assert (y.cpu() != ry).float().mean().item() < 0.01
    rel = lambda a, b: (a.float().cpu() - b.float()).abs().max().item() / max(b.float().abs().max().item(), 1e-12)  # noqa: E731
    if with_value:
        assert rel(value, rvalue) < 1e-2
        assert rel(dvw, rdvw.reshape(-1)) < 2e-2 and rel(dvb, rdvb) < 2e-2
    else:
        assert value is None and dvw is None and dvb is None
    assert rel(dx, rdx) < 2e-2 and rel(dgamma, rdg) < 2e-2


@pytest.mark.gpu
def test_final_norm_rejects_bad_shapes():
    from pipelinerl_swe_b200 import _abi, ops
    d = torch.device("cuda:0")
    with pytest.raises(_abi.B200GrpoError):
        ops.final_norm_fwd(torch.zeros(4, 12, device=d).bfloat16(), torch.ones(12, device=d).bfloat16(), 1e-6)   # H % 8
    with pytest.raises(_abi.B200GrpoError):
        ops.final_norm_fwd(torch.zeros(4, 8200, device=d).bfloat16(), torch.ones(8200, device=d).bfloat16(), 1e-6)  # H > 8192
