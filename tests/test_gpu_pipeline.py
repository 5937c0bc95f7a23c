"""The rows of the hot path chained the way the reference chains them (preprocess.py:133-170 -> finetune_loop.py:661-670,767):
traces -> reference log-probs -> prepare_rl_fields -> populate_rl_data -> packed batch -> rl_step -> step metrics."""
import copy
import os

import pytest
import torch

from helpers import StubTokenizer

pytestmark = pytest.mark.gpu


def test_traces_to_step_metrics_end_to_end():
    transformers = pytest.importorskip("transformers")
    from pipelinerl_swe_b200 import (RLConfig, StepMetrics, aggregate_rl_stats, annotate_traces_with_ref_logprobs, collate_packed,
                                     ingest_packed, populate_rl_data, prepare_rl_fields, rl_step, rl_step_async)
    d = torch.device("cuda:0")
    V = 1024 + 8
    cfg_m = transformers.Qwen2Config(vocab_size=V, hidden_size=64, intermediate_size=128, num_hidden_layers=1, num_attention_heads=4,
                                     num_key_value_heads=2, max_position_embeddings=512, tie_word_embeddings=False)
    torch.manual_seed(11)
    policy = transformers.Qwen2ForCausalLM(cfg_m).to(torch.bfloat16).to(d)
    frozen = copy.deepcopy(policy).requires_grad_(False)
    g = torch.Generator().manual_seed(2)
    eos = StubTokenizer.eos_token_id

    # 1. rollouts as the actor leaves them: token ids + behaviour log-probs of the completion (2 groups x 4 attempts)
    traces = []
    for grp in range(2):
        prompt = torch.randint(3, V, (12 + grp,), generator=g).tolist()
        for att in range(4):
            n_c = int(torch.randint(5, 40, (1,), generator=g))
            comp = torch.randint(3, V, (n_c,), generator=g).tolist() + ([eos] if att != 3 else [])
            traces.append({"input_ids": prompt + comp, "logprobs": (-torch.rand(len(comp), generator=g) * 2).tolist(),
                           "group_id": f"g{grp}", "rollout_index": att, "step_index": 0, "model_version": 1,
                           "reward": float((att + grp) % 2)})
    # 2. reference log-probs from the frozen copy (preprocess.py:144-145)
    annotate_traces_with_ref_logprobs(frozen, traces, max_tokens=256)
    # 3. + 4. RL fields and group-normalised advantages (preprocess.py:152-163)
    dataset = []
    for tr in traces:
        n, c = len(tr["input_ids"]), len(tr["logprobs"])
        enc = {"input_ids": tr["input_ids"], "labels": [-100] * (n - c) + tr["input_ids"][n - c:], "attention_mask": [1] * n}
        enc = prepare_rl_fields(enc, tr["reward"], tr["logprobs"], tr["ref_logprobs"])
        enc.update({k: tr[k] for k in ("group_id", "rollout_index", "step_index", "model_version")})
        dataset.append(enc)
    rl_cfg = RLConfig(batch_size=8, kl_coef=0.05, final_kl_coef=0.05, epsilon=0.2, divide_advantage_by_std=True)
    dataset = populate_rl_data(dataset, eos, rl_cfg)
    assert all(len(e["advantages"]) == len(e["input_ids"]) for e in dataset)
    assert dataset[3]["overflow"][0] == 1.0 and dataset[0]["overflow"][0] == 0.0        # the attempt without EOS overflowed
    # 5. two packed micro-batches through the wire form; 6. the drop-in step, with and without the per-step host sync
    halves = [dataset[:4], dataset[4:]]
    sm = StepMetrics(d)
    lists = {}
    for ex in halves:
        wire_batch = ingest_packed(copy.deepcopy(ex), StubTokenizer(), 1, d)
        ref_batch = collate_packed(copy.deepcopy(ex), StubTokenizer(), 1).to_device(d)
        loss_ref, stats = rl_step(policy, ref_batch, 0, 10, rl_cfg)
        for k, v in stats.items():
            lists.setdefault(k, []).append(v)
        loss = rl_step_async(policy, wire_batch, 0, 10, rl_cfg, sm)
        assert loss.item() == loss_ref.item()
        loss.backward()
    assert policy.lm_head.weight.grad is not None and torch.isfinite(policy.lm_head.weight.grad.float()).all()
    # 7. one reduction per optimizer step == the reference's list aggregation
    want = aggregate_rl_stats(lists, num_samples=8)
    got, counters = sm.reduce(8)
    assert counters["samples"] == 8 and counters["micro_batches"] == 2
    assert set(got) == set(want)
    for k, w in want.items():
        assert abs(got[k] - w) <= 1e-5 * max(1.0, abs(w)), (k, got[k], w)
    # the KL to the frozen copy of an unchanged policy is ~0 (ref log-probs came from the same weights)
    assert abs(got["rl/kl"]) < 1e-2


@pytest.mark.gpu
@pytest.mark.parametrize("transport", ["nccl", "symm"])
def test_vocab_parallel_two_gpus_equals_unsharded(transport):
    """tools/check_vocab_parallel.py under torchrun on 2 GPUs of this box (skipped on a 1-GPU box): NCCL transport and this
    library's NVSwitch multicast kernels (SymmetricExchange) both reproduce the unsharded log-probs / gradients."""
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541" if transport == "nccl" else "29542", os.path.join(root, "tools", "check_vocab_parallel.py")]
    if transport == "symm":
        cmd.append("--exchange=symm")
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "VOCAB_PARALLEL" in out.stdout and " OK " in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


@pytest.mark.gpu
def test_weight_update_call_shapes_and_transports_two_gpus():
    """tools/check_weight_update.py under torchrun on 2 GPUs (skipped on a 1-GPU box): a second NCCL world built by
    init_actor_update_group, WeightUpdateSender.send_weight_update(version) -> ActorUpdateWorker.receive_weight_update(request)
    over the reference loop, the bucketed NCCL path (fp32 masters, an odd-sized fp32 'direct' tensor, a ZeRO-3 stand-in) and
    the scatter + relay kernels; every received tensor bit-equal to param.data.bfloat16()."""
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29561", os.path.join(root, "tools", "check_weight_update.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "WEIGHT_UPDATE" in out.stdout and " OK " in out.stdout, out.stdout[-2000:] + out.stderr[-3000:]


def test_one_process_second_gpu_while_the_first_is_current():
    """A process that holds its model and batch on cuda:1 while cuda:0 is the current device (no ``torch.cuda.set_device``) gets
    the same step as on cuda:0 -- torch code behaves that way, so the drop-in must too: the wrappers make the tensors' device
    current for the launch, and kernel attributes (dynamic shared-memory limits) are configured per device."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from helpers import StubLM, synthetic_packed_batch
    from pipelinerl_swe_b200 import RLConfig, ops, rl_step
    T, H, V = 700, 264, 5000
    batch = synthetic_packed_batch([300, 250, 150], V, seed=4)
    cfg = RLConfig(batch_size=4, kl_coef=0.1, final_kl_coef=0.1)
    assert torch.cuda.current_device() == 0
    res = []
    for dev in ("cuda:1", "cuda:0", "cuda:1"):       # cuda:1 first: nothing has been configured on any device yet in this test
        model = StubLM(H, V, T, seed=5).to(dev)
        b = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
        loss, stats = rl_step(model, b, 0, 1, cfg)
        loss.backward()
        assert loss.device == torch.device(dev) and model.lm_head.weight.grad.device == torch.device(dev)
        res.append((loss.item(), stats, model.lm_head.weight.grad.cpu(), model.hidden.grad.cpu()))
        assert torch.cuda.current_device() == 0      # the caller's current device is left alone
    for r in res[1:]:
        assert r[0] == res[0][0] and r[1] == res[0][1]
        assert torch.equal(r[2], res[0][2]) and torch.equal(r[3], res[0][3])
    pos = torch.randint(0, 50, (5000,), device="cuda:1")
    assert ops.segment_bounds(pos).cpu().tolist() == ops.segment_bounds(pos.to("cuda:0")).cpu().tolist()
