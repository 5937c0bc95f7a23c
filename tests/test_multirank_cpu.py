"""CPU, world_size 2 over gloo: the host-side multi-rank logic (no CUDA kernels run here).

 * vocab-parallel exchange: every rank reduces its vocab shard to [M,4] partials (the oracle stands in for
   K1, which needs a GPU), ``exchange_stats`` all-gathers them, the merge equals the unsharded oracle and is
   bit-identical on both ranks (so K3 needs no further exchange).
 * weight update: sender and receivers derive the same bucket plan from one WeightUpdateRequest, and a
   bucket broadcast delivers byte-identical parameters (pack/unpack emulated with torch on CPU tensors).
"""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _init(rank, world, port):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)


def _vocab_parallel_worker(rank, world, port, q):
    _init(rank, world, port)
    import oracle
    from pipelinerl_swe_b200.vocab_parallel import exchange_stats, reduce_grad_hidden, shard_range
    g = torch.Generator().manual_seed(0)
    M, H, V, temp = 37, 16, 1003, 0.8
    hidden = torch.randn(M, H, generator=g).bfloat16()
    weight = (torch.randn(V, H, generator=g) * 0.3).bfloat16()
    ids = torch.randint(0, V, (M + 1,), generator=g)
    b, e = shard_range(V, world, rank)
    stats = oracle.oracle_shard_stats(hidden, weight[b:e], ids[1:], b, temp)
    allstats = exchange_stats(stats)
    assert allstats.shape == (world, M, 4)
    lp, ent, lse = oracle.oracle_merge_stats(allstats)
    hfull = torch.cat([hidden, hidden[:1]])
    lp_ref, ent_ref = oracle.oracle_logprobs_entropy(hfull, weight, ids, temp)
    ok = bool(torch.allclose(lp, lp_ref, rtol=1e-5, atol=1e-5) and torch.allclose(ent, ent_ref, rtol=1e-5, atol=1e-5))
    # dH = sum over shards of dlogits_r . W_r
    p = torch.exp((hidden.float() @ weight[b:e].float().t()) / temp - lse[:, None])
    gh = reduce_grad_hidden((-p) @ weight[b:e].float())
    p_full = torch.softmax((hidden.float() @ weight.float().t()) / temp, -1)
    ok = ok and bool(torch.allclose(gh, (-p_full) @ weight.float(), rtol=1e-4, atol=1e-4))
    gathered = [None] * world
    dist.all_gather_object(gathered, lp.numpy().tobytes())
    ok = ok and all(x == gathered[0] for x in gathered)  # bit-identical on every rank
    q.put((rank, ok, (b, e)))
    dist.destroy_process_group()


def _weight_update_worker(rank, world, port, q):
    _init(rank, world, port)
    from pipelinerl_swe_b200.weight_update import _bucket_elems, make_request, plan_buckets
    g = torch.Generator().manual_seed(1)
    shapes = [(64,), (33, 64), (5,), (1000, 64), (7, 3), (64,)]
    params = [(f"p{i}", torch.randn(*s, generator=g)) for i, s in enumerate(shapes)]  # fp32 master weights
    box = [make_request(params, version=3)] if rank == 0 else [None]
    dist.broadcast_object_list(box, src=0)  # stands in for the HTTP POST of finetune_loop.py:157-172
    req = box[0]
    plan = plan_buckets(req.parameters_info, 64 * 1024)
    ok = len(plan) > 1
    got = {}
    for idxs in plan:
        n = _bucket_elems(req.parameters_info, idxs)
        buf = torch.zeros(n, dtype=torch.bfloat16)
        if rank == 0:
            off = 0
            for i in idxs:
                t = params[i][1].bfloat16().reshape(-1)
                buf[off:off + t.numel()] = t
                off += (t.numel() + 7) // 8 * 8
        dist.broadcast(buf, src=0)
        off = 0
        for i in idxs:
            info = req.parameters_info[i]
            numel = 1
            for s in info.shape:
                numel *= s
            got[info.name] = buf[off:off + numel].view(tuple(info.shape)).clone()
            off += (numel + 7) // 8 * 8
    for name, p in params:  # every rank regenerates the same tensors -> byte equality with param.data.bfloat16()
        ok = ok and torch.equal(got[name].view(torch.int16), p.bfloat16().view(torch.int16))
    q.put((rank, ok, [list(b) for b in plan]))
    dist.destroy_process_group()


def _second_world_worker(rank, world, port, q):
    """A second process group beside the default one (torch_utils.py:16-65) carrying the reference-shaped update calls:
    WeightUpdateSender.send_weight_update(version) on rank 0 with a ZeRO-3 stand-in (parameters whole only inside gather),
    ActorUpdateWorker.init_actor_update_group / receive_weight_update(request) on rank 1 (vllm1.py:52-94)."""
    import contextlib
    import types
    _init(rank, world, port)                      # the process's OWN default world
    from pipelinerl_swe_b200 import weight_update as wu
    init_method = f"tcp://127.0.0.1:{port + 1}"
    g = torch.Generator().manual_seed(2)
    full = {"a.weight": torch.randn(33, 17, generator=g), "a.bias": torch.randn(17, generator=g).bfloat16(),
            "b.weight": torch.randn(5, 3, generator=g)}
    ok = True
    got = {}
    if rank != 0:   # the workers join the update group at start-up (vllm1.py:52-79), long before the first update
        class Worker(wu.ActorUpdateWorker):
            weight_update_transport = "reference"

            def __init__(self):
                self.rank, self.device = rank - 1, torch.device("cpu")
                self.model_config = types.SimpleNamespace(dtype=torch.bfloat16)
                self.model_runner = types.SimpleNamespace(model=types.SimpleNamespace(
                    load_weights=lambda weights: [got.__setitem__(n, t.clone()) or n for n, t in weights]))

        w = Worker()
        orig = wu.init_actor_update_group
        wu.init_actor_update_group = lambda *a, **k: orig(*a, **{**k, "backend": "gloo"})   # the worker mixin asks for NCCL
        try:
            w.init_actor_update_group(0, world - 1, init_method, world)
        finally:
            wu.init_actor_update_group = orig
        ok = ok and w.pg_rank == rank
    if rank == 0:
        group = wu.init_actor_update_group(init_method, 0, world, backend="gloo")
        ok = ok and dist.get_world_size(group) == world and group is not dist.group.WORLD

        class Model(torch.nn.Module):
            def __init__(self):
                super().__init__()
                self.ps = [(n, torch.nn.Parameter(t.clone(), requires_grad=False)) for n, t in full.items()]

            def named_parameters(self, *a, **k):
                return self.ps

        model = Model()
        for _, p in model.ps:                     # sharded outside gather(): half the elements, full shape in ds_shape
            p.ds_shape, p._full = p.shape, p.data
            p.data = p.data.reshape(-1)[: p.numel() // 2].clone()
        entered = []

        @contextlib.contextmanager
        def gather(params):
            entered.append(len(params))
            shards = [p.data for p in params]
            for p in params:
                p.data = p._full
            try:
                yield
            finally:
                for p, sh in zip(params, shards):
                    p.data = sh

        published = []

        def publish(req):     # the HTTP POST of finetune_loop.py:160-172, here over the default world; precedes the data
            published.append(req)
            dist.broadcast_object_list([req], src=0)
            return []

        sender = wu.WeightUpdateSender(model, group, publish=publish, transport="reference", gather=gather)
        req = sender.send_weight_update(version=11)
        ok = ok and published == [req] and req.version == 11 and [i.shape for i in req.parameters_info] == [[33, 17], [17], [5, 3]]
        ok = ok and entered == [1, 1, 1] and all(p.data.numel() == p._full.numel() // 2 for _, p in model.ps)   # re-sharded after
    else:
        box = [None]
        dist.broadcast_object_list(box, src=0)
        w.receive_weight_update(box[0])
        ok = ok and set(got) == set(full) and all(torch.equal(got[n].view(torch.int16), full[n].bfloat16().view(torch.int16)) for n in full)
    q.put((rank, ok, sorted(got)))
    dist.destroy_process_group()


def _fake_micro_batch_stats(rank, i, keys):
    """One micro-batch's stats dict as rl_step would return it (values are arbitrary but rank/step specific)."""
    g = torch.Generator().manual_seed(1000 * rank + i)
    d = {k: float(torch.randn(1, generator=g)) * 3 for k in keys}
    d["num_output_tokens_sum"] = float(100 + 7 * i + rank)
    d["input_size"] = 4096.0
    if rank == 1 and i == 1:
        d["kl"] = float("nan")            # dropped by gather_rl_metrics (finetune_loop.py:85)
        d["max_kl"] = float("inf")
    if rank == 0 and i == 2:
        d = {"input_size": 4096.0}        # micro-batch without labels (rl/__init__.py:195)
    return d


def _metrics_worker(rank, world, port, q):
    """StepMetrics' exchange + combine + finalize == the reference's all_gather_object + aggregate_rl_stats."""
    _init(rank, world, port)
    from collections import defaultdict
    from pipelinerl_swe_b200 import metrics as M
    from pipelinerl_swe_b200.stats import aggregate_rl_stats
    keys, _, kinds = M.stat_layout(has_values=True)
    K = len(keys)
    n_micro = 3 + rank                                  # ranks see different numbers of micro-batches
    # what the reference does: dict of lists per rank, pickled all-gather, finite filter, substring-keyed reduction
    lists = defaultdict(list)
    for i in range(n_micro):
        for k, v in _fake_micro_batch_stats(rank, i, keys).items():
            lists[k].append(v)
    gathered = [None] * world
    dist.all_gather_object(gathered, dict(lists))
    merged = defaultdict(list)
    for d in gathered:
        for k, v in d.items():
            merged[k].extend([x for x in v if torch.isfinite(torch.tensor(x))])
    num_samples = 11
    want = aggregate_rl_stats(merged, num_samples)
    # the accumulator the device kernel maintains, built here on the host (the kernel itself is covered on the GPU)
    acc = torch.zeros(2 * K + 4, dtype=torch.float64)
    for j, kd in enumerate(kinds):
        acc[j] = float("inf") if kd == M.KIND_MIN else float("-inf") if kd == M.KIND_MAX else 0.0
    for i in range(n_micro):
        d = _fake_micro_batch_stats(rank, i, keys)
        for j, k in enumerate(keys):
            if k in d and torch.isfinite(torch.tensor(d[k])):
                acc[j] = min(acc[j], d[k]) if kinds[j] == M.KIND_MIN else max(acc[j], d[k]) if kinds[j] == M.KIND_MAX else acc[j] + d[k]
                acc[K + j] += 1
        acc[2 * K + 1] += float(len(d) > 1)
        acc[2 * K + 2] += 1
        acc[2 * K + 3] += 5
    got, counters = M.reduce_accumulators(acc, keys, kinds, num_samples)
    ok = set(got) == set(want)
    for k in want:
        ok = ok and abs(got[k] - want[k]) <= 1e-5 * max(1.0, abs(want[k]))
    ok = ok and counters == {"labelled_micro_batches": 6, "micro_batches": 7, "samples": 35}
    # a flag raised on one rank fires the reference's assert on every rank
    acc[2 * K] = 8.0 if rank == 1 else 0.0
    try:
        M.reduce_accumulators(acc, keys, kinds, num_samples)
        ok = False
    except AssertionError as e:
        ok = ok and "Non-finite loss" in str(e)
    q.put((rank, ok, sorted(got.items())))
    dist.destroy_process_group()


def _guard(worker, rank, world, port, q):
    try:
        worker(rank, world, port, q)
    except Exception as e:  # noqa: BLE001  -- report instead of letting the parent wait for the timeout
        import traceback
        traceback.print_exc()
        q.put((rank, False, repr(e)))


def _run(worker, world=2):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_guard, args=(worker, r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
    return sorted(res)


def test_vocab_parallel_exchange_world2():
    res = _run(_vocab_parallel_worker)
    assert all(ok for _, ok, _ in res)
    (b0, e0), (b1, e1) = res[0][2], res[1][2]
    assert b0 == 0 and e0 == b1 and e1 == 1003 and (e0 - b0) % 8 == 0


def test_weight_update_plan_world2():
    res = _run(_weight_update_worker)
    assert all(ok for _, ok, _ in res)
    assert res[0][2] == res[1][2]


def test_second_process_group_and_reference_shaped_update_calls_world2():
    res = _run(_second_world_worker)
    assert res[1][2] == ["a.bias", "a.weight", "b.weight"]


def test_step_metrics_reduce_world2():
    res = _run(_metrics_worker)
    assert all(ok for _, ok, _ in res)
    assert res[0][2] == res[1][2]            # every rank ends with the same averages


def test_stat_kinds_follow_reference_substring_rule():
    from pipelinerl_swe_b200 import metrics as M
    keys, slots, kinds = M.stat_layout(True)
    by = dict(zip(keys, kinds))
    assert by["loss"] == M.KIND_SUM and by["max_loss"] == M.KIND_MAX and by["min_loss"] == M.KIND_MIN
    assert by["ratio_new_old_sum"] == M.KIND_SUM and by["num_output_tokens_sum"] == M.KIND_SUM
    assert by["value_min"] == M.KIND_MIN and by["value_max"] == M.KIND_MAX and by["value_loss"] == M.KIND_MEAN
    assert by["policy_loss"] == M.KIND_MEAN and by["kl"] == M.KIND_MEAN and by["input_size"] == M.KIND_MEAN
    assert len(keys) == 39 and len(M.stat_layout(False)[0]) == 34


def test_shard_range_covers_vocab():
    from pipelinerl_swe_b200.vocab_parallel import shard_range
    for V in (151936, 152064, 1003, 7):
        for R in (1, 2, 4, 8):
            cuts = [shard_range(V, R, r) for r in range(R)]
            assert cuts[0][0] == 0 and cuts[-1][1] == V
            assert all(a[1] == b[0] for a, b in zip(cuts[:-1], cuts[1:]))
