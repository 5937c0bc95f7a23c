"""CPU: the oracle restatement vs outputs of the reference itself (tests/golden/)."""
import copy
import json
import os

import pytest
import torch

import oracle
from helpers import GOLDEN, assert_stats_close, golden_rl_cases, load_rl_case


@pytest.mark.parametrize("name", golden_rl_cases())
def test_rl_step_matches_reference(name):
    batch, cfg, meta, x = load_rl_case(name)
    hidden = x["hidden"].clone().requires_grad_(True)
    weight = x["weight"].clone().requires_grad_(True)
    values = x["values"].clone().requires_grad_(True) if "values" in x else None
    loss, stats = oracle.oracle_rl_step(hidden, weight, batch, meta["step"], meta["max_step"], cfg, values)
    torch.testing.assert_close(loss.detach(), x["loss"], rtol=1e-5, atol=1e-6)
    assert_stats_close(stats, meta["stats"], rtol=1e-5, atol=1e-6)
    assert list(stats.keys())[:3] == ["loss", "max_loss", "min_loss"] or len(stats) == 1
    if "grad_weight" in x:
        loss.backward()
        torch.testing.assert_close(weight.grad, x["grad_weight"], rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(hidden.grad.reshape(x["grad_hidden"].shape), x["grad_hidden"], rtol=1e-4, atol=1e-6)
        if values is not None:
            torch.testing.assert_close(values.grad, x["grad_values"], rtol=1e-4, atol=1e-7)


def test_segments_bit_exact():
    pos = torch.tensor([5, 6, 7, 0, 1, 2, 0, 0, 1])
    b, n = oracle.oracle_segments(pos)
    assert b.tolist() == [0, 3, 6, 7, 9] and n == 4


ADV = json.load(open(os.path.join(GOLDEN, "advantages.json")))


@pytest.mark.parametrize("name", [k for k in ADV if k != "collate_packed"])
def test_populate_rl_data_matches_reference(name):
    case = ADV[name]
    rows = copy.deepcopy(case["input"])
    res = oracle.oracle_populate_rl_data(rows, case["eos"], oracle.OracleRLConfig(**case["cfg"]))
    assert res is rows
    for got, want in zip(res, case["output"]):
        assert got["num_labels"] == want["num_labels"]
        assert got["overflow"] == want["overflow"]
        assert got["group_tokens"] == want["group_tokens"]
        assert got["advantages"] == want["advantages"]          # f64 Python floats, bit-identical to the pandas path


def test_group_statistics_are_pandas_bit_for_bit():
    """pandas_group_mean / pandas_group_std (the restated third-party algorithm) against pandas itself: random group
    sizes and magnitudes, 0/1 rewards, singleton groups -- exact f64 equality, NaN for one sample."""
    import numpy as np
    pd = pytest.importorskip("pandas")
    rng = np.random.default_rng(0)
    for trial in range(300):
        n = int(rng.integers(1, 40))
        vals = (rng.normal(size=n) * 10.0 ** rng.integers(-3, 4) if trial % 3 == 0 else
                rng.integers(0, 2, size=n).astype(float) if trial % 3 == 1 else rng.random(n))
        r = pd.DataFrame({"g": ["a"] * n, "v": vals}).groupby("g").agg(m=("v", "mean"), s=("v", "std"))
        m, s = float(r["m"].iloc[0]), float(r["s"].iloc[0])
        assert oracle.grpo_oracle.pandas_group_mean(vals.tolist()) == m
        got = oracle.grpo_oracle.pandas_group_std(vals.tolist())
        assert got == s or (np.isnan(got) and np.isnan(s))


def test_collate_packed_matches_reference():
    case = ADV["collate_packed"]
    got = oracle.oracle_collate_packed(copy.deepcopy(case["examples"]), case["eos"], case["seq_parallel"])
    for k, want in case["batch"].items():
        g = got[k]
        if isinstance(g, torch.Tensor):
            assert g.tolist() == want, k
            if k == "seq_boundaries":
                assert g.dtype == torch.int32
        else:
            assert g == want, k


def test_aggregate_matches_reference():
    case = json.load(open(os.path.join(GOLDEN, "aggregate.json")))
    got = oracle.oracle_aggregate_rl_stats(case["input"], case["num_samples"])
    assert got.keys() == case["output"].keys()
    for k in got:
        assert abs(got[k] - case["output"][k]) < 1e-6


def test_filter_zero_advantage_groups():
    rows = [dict(group_id="a", advantages=[0.0, 1e-7]), dict(group_id="b", advantages=[0.0, 0.5]), dict(group_id="a", advantages=[0.0])]
    kept, dropped = oracle.oracle_filter_zero_advantage_groups(rows)
    assert dropped == 2 and [r["group_id"] for r in kept] == ["b"]


def test_sentinel_batch_layout():
    sb = oracle.oracle_sentinel_batch(2, 7)
    import numpy as np
    z = np.load(os.path.join(GOLDEN, "rl_step_sentinel.npz"))
    for k in ("input_ids", "labels", "position_ids", "group_tokens", "num_labels", "seq_boundaries"):
        assert sb[k].tolist() == z[k].tolist(), k


@pytest.mark.parametrize("case", ["a", "b"])
def test_final_norm_value_matches_transformers_and_reference_value_head(case):
    """oracle_final_norm_value vs transformers' Qwen2RMSNorm + the reference's ValueHead (bf16, with autograd)."""
    import numpy as np
    z = np.load(os.path.join(GOLDEN, "final_norm.npz"))
    t = lambda k: torch.from_numpy(z[f"{case}_{k}"]).to(torch.bfloat16)  # noqa: E731
    x, gamma, vw, vb = t("x").requires_grad_(True), t("gamma").requires_grad_(True), t("vw").requires_grad_(True), t("vb").requires_grad_(True)
    y, value = oracle.oracle_final_norm_value(x, gamma, 1e-6, vw, vb)
    assert torch.equal(y.detach(), t("y")) and torch.equal(value.detach(), t("value"))
    torch.autograd.backward([y, value], [t("dy"), t("dv")])
    assert torch.equal(x.grad, t("dx")) and torch.equal(gamma.grad, t("dgamma"))
    assert torch.equal(vw.grad, t("dvw")) and torch.equal(vb.grad, t("dvb"))
