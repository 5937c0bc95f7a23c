"""CPU: host logic of the drop-in layer vs the reference's golden outputs, the C-ABI library's exported
symbols vs include/b200grpo.h, and loud failure without a GPU (no CPU fallback)."""
import copy
import ctypes
import json
import os
import re

import numpy as np
import pytest
import torch

import oracle
from helpers import GOLDEN, StubLM, synthetic_packed_batch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ADV = json.load(open(os.path.join(GOLDEN, "advantages.json")))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "b200grpo.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200grpo_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_header_symbol():
    from pipelinerl_swe_b200 import _abi
    names = _header_symbols()
    assert len(names) >= 20
    handle = ctypes.CDLL(_abi.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), f"libb200grpo.so does not export {n}"
    assert set(names) == set(_abi.SIGNATURES), set(names) ^ set(_abi.SIGNATURES)
    assert _abi.lib().b200grpo_abi_version() == 3


def test_struct_layouts_match_header():
    from pipelinerl_swe_b200 import _abi
    assert ctypes.sizeof(_abi.LossCfg) == 8 * 4 + 6 * 4
    assert ctypes.sizeof(_abi.HostBatch) == 11 * ctypes.sizeof(ctypes.c_void_p)
    src = open(os.path.join(ROOT, "include", "b200grpo.h")).read()
    body = src[src.index("B200GRPO_M_LOSS = 0"):src.index("B200GRPO_M_COUNT")]
    slots = re.findall(r"B200GRPO_(M_[A-Z0-9_]+)", body)
    for i, s in enumerate(slots):
        assert getattr(_abi, s) == i, s


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    from pipelinerl_swe_b200 import RLConfig, _abi, ops, rl_step
    batch = synthetic_packed_batch([20, 12], 100, seed=2)
    with pytest.raises(_abi.B200GrpoError):
        rl_step(StubLM(32, 100, 32), batch, 0, 1, RLConfig(batch_size=2))
    with pytest.raises(_abi.B200GrpoError):
        ops.segment_bounds(torch.zeros(8, dtype=torch.int64))
    # the C side refuses too: no device -> error code + message, never a computed answer
    rc = _abi.lib().b200grpo_lmhead_finalize(None, 1, 1, None, None, None, None)
    assert rc != 0 and _abi.lib().b200grpo_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "pipelinerl_swe_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f


# ---- host-side mirrors of the reference's data model, vs golden ---------------------------------
def test_collate_packed_vs_reference_golden():
    from pipelinerl_swe_b200 import collate_packed
    case = ADV["collate_packed"]
    tok = type("Tok", (), {"eos_token_id": case["eos"]})()
    got = collate_packed(copy.deepcopy(case["examples"]), tok, case["seq_parallel"])
    for k, want in case["batch"].items():
        g = getattr(got, k)
        if isinstance(g, torch.Tensor):
            assert g.tolist() == want, k
        else:
            assert g == want, k
    assert got.seq_boundaries.dtype == torch.int32 and got.input_ids.dtype == torch.int64
    assert got.rewards.dtype == torch.float32


def test_make_slices_and_sentinel():
    from pipelinerl_swe_b200 import collate_packed, create_sentinel_batch
    case = ADV["collate_packed"]
    tok = type("Tok", (), {"eos_token_id": 2})()
    pb = collate_packed(copy.deepcopy(case["examples"]), tok, 4)
    sl = pb.make_slices(4)
    assert len(sl) == 4 and all(s.input_ids.shape[1] == pb.input_ids.shape[1] // 4 for s in sl)
    assert torch.equal(torch.cat([s.labels for s in sl], 1), pb.labels)
    with pytest.raises(ValueError):
        pb.make_slices(7)
    sb = create_sentinel_batch(None, tok, model_version=7)
    z = np.load(os.path.join(GOLDEN, "rl_step_sentinel.npz"))
    for k in ("input_ids", "labels", "position_ids", "group_tokens", "num_labels", "seq_boundaries"):
        assert getattr(sb, k).tolist() == z[k].tolist(), k
    assert sb.sentinel and sb.is_packed and sb.model_version == 7


def test_prepare_rl_fields_and_filter():
    from pipelinerl_swe_b200 import filter_zero_advantage_groups, prepare_rl_fields
    enc = dict(input_ids=[5, 6, 7, 8], labels=[-100, -100, 7, 8])
    got = prepare_rl_fields(dict(enc), 0.5, [-0.1, -0.2], [-0.3, -0.4])
    want = oracle.oracle_prepare_rl_fields(dict(enc), 0.5, [-0.1, -0.2], [-0.3, -0.4])
    assert got == want
    with pytest.raises(AssertionError):
        prepare_rl_fields(dict(enc), 0.5, [-0.1], [-0.3])
    rows = [dict(group_id="a", advantages=[0.0, 1e-7]), dict(group_id="b", advantages=[0.0, 0.5]), dict(group_id="a", advantages=[0.0])]
    kept, dropped = filter_zero_advantage_groups(rows)
    assert dropped == 2 and [r["group_id"] for r in kept] == ["b"]


def test_aggregate_rl_stats_vs_reference_golden():
    from pipelinerl_swe_b200 import aggregate_rl_stats
    case = json.load(open(os.path.join(GOLDEN, "aggregate.json")))
    got = aggregate_rl_stats(case["input"], case["num_samples"])
    assert got.keys() == case["output"].keys()
    for k in got:
        assert abs(got[k] - case["output"][k]) < 1e-6


def test_rlconfig_ignores_unknown_keys_like_reference():
    from pipelinerl_swe_b200 import RLConfig, linear_decay_coef
    c = RLConfig(policy_loss="reinforce", aggregate_loss="sum", reward_minus_kl_coef=0.3)
    assert c.policy_loss == "reinforce" and not hasattr(c, "aggregate_loss")
    assert linear_decay_coef(3, 10, 0.1, 0.2) == oracle.oracle_linear_decay_coef(3, 10, 0.1, 0.2)


def test_rollout_scalars_host_side():
    from pipelinerl_swe_b200.advantage import rollout_scalars
    case = ADV["singleton_and_steps"]
    reward, gkey, offsets, nlab, eos, n_groups, tok, order, starts = rollout_scalars(copy.deepcopy(case["input"]), case["eos"])
    rows = case["input"]
    assert offsets.tolist() == np.concatenate([[0], np.cumsum([len(r["input_ids"]) for r in rows])]).tolist()
    assert n_groups == len({f"{r['group_id']}_{r['step_index']}" for r in rows})
    assert nlab.tolist() == [sum(t != -100 for t in r["labels"]) for r in rows]
    assert eos.tolist() == [int(case["eos"] in r["input_ids"]) for r in rows]
    assert tok is None and reward.dtype == np.float64
    # the CSR K4 walks: rows stably sorted by group key (dataset order inside a group), group g at order[starts[g]:starts[g+1]]
    assert order.dtype == np.int32 and starts.dtype == np.int32 and starts[0] == 0 and starts[-1] == len(rows)
    for g in range(n_groups):
        members = order[starts[g]:starts[g + 1]].tolist()
        assert members == [i for i in range(len(rows)) if gkey[i] == g] and members == sorted(members)


def test_plan_buckets():
    from pipelinerl_swe_b200.weight_update import ParameterInfo, plan_buckets
    infos = [ParameterInfo(name=f"p{i}", shape=s, dtype="torch.bfloat16") for i, s in
             enumerate([[3584], [3584, 3584], [5], [152064, 3584], [18944, 3584], [7]])]
    plan = plan_buckets(infos, 64 << 20)
    assert [i for b in plan for i in b] == list(range(len(infos)))  # order preserved, nothing lost
    assert [3] in plan                                               # oversize parameter gets its own bucket
    cap = (64 << 20) // 2
    for b in plan:
        n = sum((int(np.prod(infos[i].shape)) + 7) // 8 * 8 for i in b)
        assert n <= cap or len(b) == 1
    assert plan_buckets(infos, 1 << 40) == [list(range(len(infos)))]


def test_plan_transfers_covers_every_parameter_once():
    from pipelinerl_swe_b200.weight_update import ParameterInfo, plan_transfers
    infos = [ParameterInfo(name=f"p{i}", shape=s, dtype="torch.bfloat16") for i, s in
             enumerate([[3584], [3584, 3584], [512, 3584], [152064, 3584], [18944, 3584], [7], [512]])]
    plan = plan_transfers(infos, 64 << 20, direct_bytes=8 << 20)
    seen = []
    for kind, item in plan:
        seen += item if kind == "bucket" else [item]
    assert sorted(seen) == list(range(len(infos)))
    assert ("direct", 1) in plan and ("direct", 3) in plan and ("direct", 4) in plan
    assert [k for k, _ in plan].count("bucket") == 1 and plan[0] == ("bucket", [0, 2, 5, 6])


def test_ref_logprob_pack_plan():
    from pipelinerl_swe_b200.ref_logprobs import completion_spans, plan_packs
    assert plan_packs([30, 200, 64, 150, 90, 400], 300) == [[0, 1, 2], [3, 4], [5]]
    assert plan_packs([300, 1, 299], 300) == [[0], [1, 2]]
    assert plan_packs([], 10) == []
    with pytest.raises(ValueError):
        plan_packs([3, 0], 10)
    assert completion_spans([{"input_ids": [1, 2, 3], "logprobs": [0.1]}]) == [(3, 1)]
    with pytest.raises(ValueError):
        completion_spans([{"input_ids": [1, 2], "logprobs": [0.1, 0.2]}])


def test_packed_wire_layout_matches_collate_packed():
    """Host side of the 8f-1 ingest: the wire sections hold what collate_packed would put in its tensors."""
    from helpers import StubTokenizer, synthetic_rollout_examples
    from pipelinerl_swe_b200 import collate_packed
    from pipelinerl_swe_b200.ingest import SEQ_COLUMNS, NonUniformColumns, pack_rollouts
    ex = synthetic_rollout_examples([5, 1, 9, 0, 4], 100, seed=3)
    for sp in (1, 8):
        want = collate_packed(copy.deepcopy(ex), StubTokenizer(), sp)
        wire = pack_rollouts(copy.deepcopy(ex), StubTokenizer(), sp, pin=False)
        T = want.input_ids.shape[1]
        assert wire.T == T and wire.padding == want.padding and wire.model_version == want.model_version
        assert wire.n_seq == len(want.seq_boundaries) - 1
        assert torch.equal(wire.section("input_ids").long(), want.input_ids[0])
        assert torch.equal(wire.section("labels").long(), want.labels[0])
        assert torch.equal(wire.section("old_logprobs"), want.old_logprobs[0])
        assert torch.equal(wire.section("ref_logprobs"), want.ref_logprobs[0])
        assert torch.equal(wire.section("seq_boundaries"), want.seq_boundaries)
        scal = wire.section("seq_scalars").reshape(5, wire.n_seq)
        b = want.seq_boundaries.tolist()
        for r, col in enumerate(SEQ_COLUMNS):
            for s in range(wire.n_seq):
                if b[s + 1] > b[s]:
                    assert scal[r, s] == getattr(want, col)[0, b[s]]
        assert wire.nbytes <= 16 * T + 24 * wire.n_seq + 6 * 16 + 4
    with pytest.raises(NonUniformColumns):
        pack_rollouts(synthetic_rollout_examples([5, 7], 100, per_token_rewards=True), StubTokenizer(), 1, pin=False)


def test_install_fused_lm_head_keeps_module_identity_and_hooks():
    """The swap is an in-place re-class: same module object, same Parameter, hooks registered earlier still fire."""
    from pipelinerl_swe_b200 import FusedLMHead, install_fused_lm_head

    class Wrapper(torch.nn.Module):      # DeepSpeed-engine-like nesting
        def __init__(self):
            super().__init__()
            self.module = torch.nn.Module()
            self.module.lm_head = torch.nn.Linear(16, 40, bias=False)

    w = Wrapper()
    lin = w.module.lm_head
    calls = []
    lin.register_forward_pre_hook(lambda m, a: calls.append("pre"))
    keys = list(w.state_dict().keys())
    head = install_fused_lm_head(w)
    assert head is lin and w.module.lm_head is lin and isinstance(lin, FusedLMHead) and isinstance(lin, torch.nn.Linear)
    assert install_fused_lm_head(w) is lin and type(lin).__mro__.count(FusedLMHead) == 1   # idempotent
    assert list(w.state_dict().keys()) == keys
    x = torch.randn(2, 3, 16)
    assert torch.equal(lin(x), torch.nn.functional.linear(x, lin.weight)) and calls == ["pre"]   # plain projection outside rl_step
    with pytest.raises(ValueError):
        install_fused_lm_head(torch.nn.ModuleDict({"lm_head": torch.nn.Linear(4, 4, bias=True)}))
    standalone = FusedLMHead(torch.nn.Linear(16, 40, bias=False))
    assert standalone(x).shape == (2, 3, 40)


def test_install_fused_final_norm_is_an_in_place_reclass():
    """Same module objects and state dict; CPU tensors keep the module's own forward (the kernel has no CPU path); the
    value head falls back to its own Linear when it is not handed the fused norm's output; unsuitable norms are refused."""
    from pipelinerl_swe_b200 import install_fused_final_norm
    from pipelinerl_swe_b200.final_norm import FusedFinalNorm, FusedValueHead

    class RMSNorm(torch.nn.Module):            # the transformers Qwen2RMSNorm / LlamaRMSNorm shape
        def __init__(self, H, eps=1e-6):
            super().__init__()
            self.weight = torch.nn.Parameter(torch.ones(H))
            self.variance_epsilon = eps

        def forward(self, x):
            h = x.float()
            h = h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + self.variance_epsilon)
            return self.weight * h.to(x.dtype)

    class Decoder(torch.nn.Module):
        def __init__(self, H):
            super().__init__()
            self.layers = torch.nn.ModuleList([torch.nn.Linear(H, H)])
            self.norm = RMSNorm(H)

    class LM(torch.nn.Module):
        def __init__(self, H):
            super().__init__()
            self.model = Decoder(H)
            self.lm_head = torch.nn.Linear(H, 32, bias=False)

    class ValueHead(torch.nn.Module):          # pipelinerl/finetune/value_model.py:40-52
        def __init__(self, H):
            super().__init__()
            self.output = torch.nn.Linear(H, 1)

        def forward(self, hidden_states):
            return self.output(hidden_states).squeeze(-1)

    class WithValueHead(torch.nn.Module):      # value_model.py:55-66
        def __init__(self, lm, H):
            super().__init__()
            self.pretrained_model = lm
            self.value_head = ValueHead(H)

    m = WithValueHead(LM(16), 16)
    norm, head = m.pretrained_model.model.norm, m.value_head
    keys = list(m.state_dict().keys())
    got = install_fused_final_norm(m)
    assert got is norm and isinstance(norm, FusedFinalNorm) and isinstance(norm, RMSNorm) and isinstance(head, FusedValueHead)
    assert install_fused_final_norm(m) is norm and type(norm).__mro__.count(FusedFinalNorm) == 1
    assert list(m.state_dict().keys()) == keys and "_b200_value_linear" not in dict(norm.named_modules())
    x = torch.randn(3, 5, 16)
    assert torch.equal(norm(x), RMSNorm.forward(norm, x))                 # CPU tensors: the module's own forward
    assert torch.equal(head(x), head.output(x).squeeze(-1))              # not the fused norm's output: own Linear
    bad = LM(16)
    del bad.model.norm.variance_epsilon
    with pytest.raises(ValueError):
        install_fused_final_norm(bad)
    with pytest.raises(AttributeError):
        install_fused_final_norm(torch.nn.Linear(4, 4))


def test_row_chunks_respect_the_budget():
    """set_saved_logits_budget: every chunk's rows x padded vocab x 2 B fits the budget (floor 4096 rows), chunks tile the
    rows exactly, are multiples of 512 rows, balanced, and a micro-batch that fits is never cut."""
    from pipelinerl_swe_b200.lm_head import row_chunks
    V = 152064
    per_row = -(-V // 256) * 256 * 2
    assert row_chunks(8191, V) == [(0, 8191)]
    for M, budget in ((20000, 4 << 30), (20000, 2 << 30), (49999, 4 << 30), (50000, 1 << 28)):
        ch = row_chunks(M, V, budget)
        assert ch[0][0] == 0 and ch[-1][1] == M and all(a[1] == b[0] for a, b in zip(ch, ch[1:]))
        fit = max(4096, budget // per_row // 512 * 512)
        assert all(r1 - r0 <= fit for r0, r1 in ch) and all((r1 - r0) % 512 == 0 for r0, r1 in ch[:-1])
        assert len(ch) == -(-M // fit)                                  # as few chunks as the budget allows
        assert ch[-1][1] - ch[-1][0] > M - (len(ch) - 1) * fit - 1          # and balanced: the last one is not the sliver M % fit
    assert row_chunks(49999, 19008) == [(0, 49999)]   # a vocab shard of eight: 50k rows fit the default budget


def test_install_with_vocab_group_records_the_shard():
    """install_fused_lm_head(model, vocab_group=..., v_offset=...) marks the head as a vocab shard; rl_step's request picks it up."""
    from pipelinerl_swe_b200 import install_fused_lm_head
    m = torch.nn.ModuleDict({"lm_head": torch.nn.Linear(16, 40, bias=False)})
    grp = object()
    head = install_fused_lm_head(m, vocab_group=grp, v_offset=80)
    assert head.vocab_group is grp and head.v_offset == 80
    assert install_fused_lm_head(m) is head and head.vocab_group is grp          # a later plain install keeps it
    assert "vocab_group" not in dict(head.named_parameters()) and list(m.state_dict().keys()) == ["lm_head.weight"]


def test_rl_step_signature_is_the_reference_plus_keyword_only_extras():
    import inspect
    from pipelinerl_swe_b200 import rl_step
    params = list(inspect.signature(rl_step).parameters.values())
    assert [p.name for p in params[:5]] == ["model", "batch", "current_step", "max_step", "config"]   # rl/__init__.py:130-136
    assert all(p.kind is inspect.Parameter.KEYWORD_ONLY and p.default is None for p in params[5:])


def test_staged_reference_reproduces_the_golden_vectors_and_the_oracle():
    """oracle/_ref (the reference's own files staged by oracle/build_ref.py; what bench.py's CPU arm runs) gives the committed
    golden loss / stats, i.e. it IS the code that pinned the oracle.  Skipped where neither _ref nor /root/reference exists."""
    import types
    from oracle import build_ref
    if not build_ref.available() and not build_ref.build():
        pytest.skip("oracle/_ref not staged and /root/reference absent")
    rl = build_ref.load()
    import sys
    PBE = sys.modules["pipelinerl.finetune.types"].PipelineBatchEncoding
    from helpers import load_rl_case
    import oracle
    for name in ("ppo_kl_packed", "reinforce_gn_of_ent"):
        batch, ocfg, meta, x = load_rl_case(name)
        h = x["hidden"].float().requires_grad_(True)
        w = x["weight"].float().requires_grad_(True)
        hh = h if h.dim() == 3 else h[None]

        class M(torch.nn.Module):
            def forward(self, **kw):
                return types.SimpleNamespace(logits=hh @ w.t())

        b = {k: v for k, v in batch.items() if k != "seq_boundaries"}
        loss, stats = rl.rl_step(M(), PBE(**b, model_version=0), meta["step"], meta["max_step"], rl.RLConfig(**meta["cfg"]))
        assert abs(float(loss) - float(x["loss"])) <= 1e-6 * abs(float(x["loss"])) + 1e-7
        for k, v in meta["stats"].items():
            assert abs(stats[k] - v) <= 1e-5 * abs(v) + 1e-6, (name, k)
        o_loss, _ = oracle.oracle_rl_step(x["hidden"].float(), x["weight"].float(), batch, meta["step"], meta["max_step"], ocfg)
        assert abs(float(o_loss) - float(loss)) <= 1e-5 * abs(float(loss)) + 1e-6


def test_make_slices_can_carry_the_boundary_target():
    """Default = the reference's cuts (types.py:144-178).  carry_boundary_target=True attaches the next slice's first token to
    every slice but the last, so that the slices together hold every (position, next token) pair of the un-sliced row."""
    from pipelinerl_swe_b200 import PipelineBatchEncoding
    from helpers import synthetic_packed_batch
    b = synthetic_packed_batch([10, 14, 8], 50, seed=1)
    pbe = PipelineBatchEncoding(**b, model_version=0)
    plain = pbe.make_slices(4)
    carried = pbe.make_slices(4, carry_boundary_target=True)
    assert all(s.boundary is None for s in plain) and carried[-1].boundary is None
    pairs = []
    for k, s in enumerate(carried):
        assert torch.equal(s.input_ids, plain[k].input_ids) and s.input_ids.shape == (1, 8)     # the cuts themselves do not move
        ids = s.input_ids[0].tolist()
        pairs += list(zip(ids[:-1], ids[1:]))
        if s.boundary is not None:
            assert set(s.boundary) >= {"input_ids", "labels", "rewards", "advantages", "old_logprobs", "ref_logprobs", "position_ids"}
            assert all(v.shape == (1, 1) for v in s.boundary.values())
            assert s.boundary["input_ids"][0, 0] == carried[k + 1].input_ids[0, 0]
            assert s.boundary["old_logprobs"][0, 0] == carried[k + 1].old_logprobs[0, 0]
            pairs.append((ids[-1], int(s.boundary["input_ids"][0, 0])))
    full = pbe.input_ids[0].tolist()
    assert pairs == list(zip(full[:-1], full[1:]))


def test_batch_type_behaves_like_the_reference_batch_type():
    """PipelineBatchEncoding (validators, make_slices, to_device, error messages) against the reference's own class staged in
    oracle/_ref (finetune/types.py:48-178), on random packed rows built from lists / ndarrays / tensors of the wrong dtype."""
    import random

    import numpy as np
    from oracle import build_ref
    if not build_ref.available():
        pytest.skip("oracle/_ref is not staged on this machine")
    build_ref.load()
    import sys as _sys
    RefPBE = _sys.modules["pipelinerl.finetune.types"].PipelineBatchEncoding
    from pipelinerl_swe_b200.types import PipelineBatchEncoding as OurPBE
    long_f = ("input_ids", "attention_mask", "labels", "position_ids")
    float_f = ("rewards", "advantages", "ref_logprobs", "old_logprobs", "group_tokens", "num_labels", "overflow")
    rnd = random.Random(5)
    for case in range(24):
        n_slices = rnd.choice([1, 2, 4, 8])
        T = n_slices * rnd.randint(1, 40)
        b = synthetic_packed_batch([T], 500, seed=case)
        form = rnd.choice(["tensor", "list", "ndarray", "wrong_dtype"])
        kw = {}
        for k in long_f + float_f:
            v = b[k]
            if form == "list":
                v = v.tolist()
            elif form == "ndarray":
                v = v.numpy()
            elif form == "wrong_dtype":
                v = v.double() if k in float_f else v.int()
            kw[k] = v
        kw.update(model_version=case, is_packed=True, seq_boundaries=[0, T], padding=rnd.randint(0, 3), sentinel=bool(case % 5 == 0))
        ours, ref = OurPBE(**kw), RefPBE(**kw)
        for k in long_f + float_f + ("seq_boundaries",):
            a, r = getattr(ours, k), getattr(ref, k)
            assert a.dtype == r.dtype and a.shape == r.shape and torch.equal(a, r), k
        so, sr = ours.make_slices(n_slices), ref.make_slices(n_slices)
        assert len(so) == len(sr) == n_slices
        for x, y in zip(so, sr):
            for k in long_f + float_f:
                assert torch.equal(getattr(x, k), getattr(y, k)), k
            for k in ("model_version", "sentinel", "is_packed", "padding"):
                assert getattr(x, k) == getattr(y, k), k
            assert torch.equal(x.seq_boundaries, y.seq_boundaries)
            assert getattr(x, "boundary", None) is None          # the carried boundary target is opt-in
        assert ours.to_device("cpu") is ours
    # the three ways make_slices refuses, with the reference's exception type and text
    b = synthetic_packed_batch([12], 50, seed=1)
    base = {k: b[k] for k in long_f + float_f}
    for mutate, n in ((lambda d: d.update(position_ids=None), 2), (lambda d: None, 13), (lambda d: None, 5)):
        d = dict(base, model_version=0, is_packed=True)
        mutate(d)
        msgs = []
        for cls in (OurPBE, RefPBE):
            with pytest.raises(ValueError) as ei:
                cls(**d).make_slices(n)
            msgs.append(str(ei.value))
        assert msgs[0] == msgs[1], msgs
    for cls in (OurPBE, RefPBE):     # unsupported input type
        with pytest.raises(Exception):
            cls(**dict(base, model_version=0, input_ids="abc"))
