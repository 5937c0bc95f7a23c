"""GPU parity: the CUDA path (through the C ABI) against the oracle and the reference's golden vectors.

Tolerances are the north-star's: logprobs / loss 1e-3 relative (bf16 inputs, fp32 accumulate),
gradients 2e-2 relative (max-norm), advantages and segment indexing bit-exact.
"""
import copy
import json
import os

import numpy as np
import pytest
import torch

import oracle
from helpers import (GOLDEN, StubLM, assert_stats_close, golden_rl_cases, load_rl_case, near_policy_logprobs,
                     synthetic_packed_batch)

pytestmark = pytest.mark.gpu

LP_RTOL = 1e-3   # logprobs, loss, stats
GRAD_RTOL = 2e-2  # gradients, relative to the largest reference entry


def _dev():
    return torch.device("cuda:0")


def _to_dev(batch):
    return {k: (v.to(_dev()) if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}


def _rel(a, b):
    return float((a.float().cpu() - b.float().cpu()).abs().max() / (b.float().abs().max() + 1e-20))


class GoldenLM(torch.nn.Module):
    """hidden/weight of a golden case as a model whose forward goes through lm_head."""

    def __init__(self, hidden, weight, values=None):
        super().__init__()
        if hidden.dim() == 2:
            hidden = hidden[None]
        self.hidden = torch.nn.Parameter(hidden.to(torch.bfloat16))
        V, H = weight.shape
        self.lm_head = torch.nn.Linear(H, V, bias=False, dtype=torch.bfloat16)
        with torch.no_grad():
            self.lm_head.weight.copy_(weight.to(torch.bfloat16))
        if values is not None:
            self.value_head = True
            self.values = torch.nn.Parameter(values.float())

    def forward(self, **kw):
        import types
        out = types.SimpleNamespace(logits=self.lm_head(self.hidden))
        if hasattr(self, "values"):
            out.value = self.values
        return out


# ---------------------------------------------------------------------------------------------
# rl_step end to end vs the REFERENCE's own outputs (tests/golden, generated from /root/reference)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", golden_rl_cases())
def test_rl_step_vs_reference_golden(name):
    from pipelinerl_swe_b200 import RLConfig, rl_step
    batch, ocfg, meta, x = load_rl_case(name)
    model = GoldenLM(x["hidden"], x["weight"], x.get("values")).to(_dev())
    cfg = RLConfig(**meta["cfg"])
    loss, stats = rl_step(model, _to_dev(batch), meta["step"], meta["max_step"], cfg)
    assert loss.dim() == 0 and loss.requires_grad
    want = float(x["loss"])
    assert abs(loss.item() - want) <= LP_RTOL * abs(want) + 1e-5, (loss.item(), want)
    assert_stats_close(stats, meta["stats"], rtol=2e-3, atol=2e-4)
    assert list(stats.keys()) == list(meta["stats"].keys()) or set(stats) == set(meta["stats"])
    if "grad_weight" in x:
        loss.backward()
        assert _rel(model.lm_head.weight.grad, x["grad_weight"]) < GRAD_RTOL
        assert _rel(model.hidden.grad.reshape(x["grad_hidden"].shape), x["grad_hidden"]) < GRAD_RTOL
        if "grad_values" in x:
            assert _rel(model.values.grad, x["grad_values"]) < GRAD_RTOL
    else:  # sentinel: loss*0 must still backprop (finetune_loop.py:664-666)
        (loss * 0.0).backward()
        assert torch.count_nonzero(model.lm_head.weight.grad) == 0


# ---------------------------------------------------------------------------------------------
# K1 / K2 at the real vocab size vs the oracle (fp32 math on the same bf16 values)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("save", [False, True, "no_entropy"])
@pytest.mark.parametrize("two_cta", [0, 1])
@pytest.mark.parametrize("T,H,V,temp", [(513, 896, 151936, 1.0), (300, 1536, 152064, 0.7), (260, 256, 1000 + 8, 1.3)])
def test_lmhead_fwd_bwd_vs_oracle(T, H, V, temp, two_cta, save):
    """save=False: backward recomputes the logits; save=True: forward keeps bf16 exponentials (relative to a per-row reference)
    and the backward turns them into dlogits in place; "no_entropy": no gradient through the entropy -> the backward's GEMMs
    read the saved exponentials directly (row-scale epilogue, scaled hidden, sparse target rows; bf16 atomic path of dW)."""
    from pipelinerl_swe_b200 import _abi, ops
    _abi.set_tunable("two_cta", two_cta)
    try:
        g = torch.Generator().manual_seed(T + H)
        hidden = torch.randn(T, H, generator=g).bfloat16()
        weight = (torch.randn(V, H, generator=g) * 0.03).bfloat16()
        ids = torch.randint(0, V, (T,), generator=g)
        hr, wr = hidden.float().requires_grad_(True), weight.float().requires_grad_(True)
        lp_ref, ent_ref = oracle.oracle_logprobs_entropy(hr, wr, ids, temp)
        c_lp = torch.randn(T - 1, generator=g) * 0.1
        c_lp[::7] = 0.0                                                     # unlabelled positions carry a zero coefficient
        c_ent = torch.randn(T - 1, generator=g) * 0.01 * (save != "no_entropy")
        ((lp_ref * c_lp).sum() + (ent_ref * c_ent).sum()).backward()

        d = _dev()
        h, w, tgt = hidden[:-1].contiguous().to(d), weight.to(d), ids[1:].contiguous().to(d)
        saved = ops.lmhead_saved_buffer(T - 1, H, V, d) if save else None
        stats = ops.lmhead_fwd_stats(h, w, tgt, 1.0 / temp, saved=saved)
        lp, ent, lse = ops.lmhead_finalize(stats)
        assert float((lp.cpu() - lp_ref.detach()).abs().max()) <= LP_RTOL * float(lp_ref.detach().abs().max())
        assert float((ent.cpu() - ent_ref.detach()).abs().max()) <= LP_RTOL * float(ent_ref.detach().abs().max())
        gh, gw = ops.lmhead_bwd(h, w, tgt, lse, ent, c_lp.to(d), None if save == "no_entropy" else c_ent.to(d), 1.0 / temp, saved=saved)
        assert _rel(gh, hr.grad[:-1]) < GRAD_RTOL
        assert _rel(gw, wr.grad) < GRAD_RTOL
        if save == "no_entropy":
            assert torch.count_nonzero(gh[::7]) == 0                       # zero coefficient -> exactly zero gradient row
    finally:
        _abi.set_tunable("two_cta", 1)


def test_lmhead_vocab_shards_equal_unsharded():
    """Config-4 semantics on one GPU: R contiguous vocab shards (v_offset) merged by finalize == unsharded."""
    from pipelinerl_swe_b200 import ops
    d = _dev()
    g = torch.Generator(device=d).manual_seed(4)
    M, H, V = 700, 512, 152064
    h = torch.randn(M, H, device=d, generator=g).bfloat16()
    w = (torch.randn(V, H, device=d, generator=g) * 0.04).bfloat16()
    tgt = torch.randint(0, V, (M,), device=d, generator=g)
    lp0, ent0, lse0 = ops.lmhead_finalize(ops.lmhead_fwd_stats(h, w, tgt, 1.0))
    c_lp = torch.randn(M, device=d, generator=g) * 0.1
    c_ent = torch.randn(M, device=d, generator=g) * 0.01
    gh0, gw0 = ops.lmhead_bwd(h, w, tgt, lse0, ent0, c_lp, c_ent, 1.0)
    for R in (2, 4, 8):
        Vs = V // R
        parts = [ops.lmhead_fwd_stats(h, w[r * Vs:(r + 1) * Vs], tgt, 1.0, v_offset=r * Vs) for r in range(R)]
        lp, ent, lse = ops.lmhead_finalize(torch.stack(parts))
        assert float((lp - lp0).abs().max()) < 1e-4 and float((ent - ent0).abs().max()) < 1e-4
        gh = torch.zeros(M, H, device=d)
        for r in range(R):
            gh_r, gw_r = ops.lmhead_bwd(h, w[r * Vs:(r + 1) * Vs], tgt, lse, ent, c_lp, c_ent, 1.0, v_offset=r * Vs)
            gh += gh_r.float()
            assert _rel(gw_r, gw0[r * Vs:(r + 1) * Vs]) < 1e-2
        assert _rel(gh, gh0) < 1e-2


def test_lmhead_properties_full_size():
    """BASELINE config-2 size (T=8192, H=1536, V=151936): size-independent properties."""
    from pipelinerl_swe_b200 import ops
    d = _dev()
    g = torch.Generator(device=d).manual_seed(9)
    M, H, V = 8191, 1536, 151936
    h = torch.randn(M, H, device=d, generator=g).bfloat16()
    w = (torch.randn(V, H, device=d, generator=g) * 0.02).bfloat16()
    tgt = torch.randint(0, V, (M,), device=d, generator=g)
    lp, ent, lse = ops.lmhead_finalize(ops.lmhead_fwd_stats(h, w, tgt, 1.0))
    assert bool((lp <= 0).all()) and bool(torch.isfinite(lp).all())
    assert bool((ent >= 0).all()) and float(ent.max()) <= float(np.log(V)) + 1e-3
    # spot rows against a torch fp32 computation of the same rows
    rows = torch.tensor([0, 1, 127, 128, 4095, 8190], device=d)
    z = h[rows].float() @ w.float().t()
    logp = torch.log_softmax(z, -1)
    assert float((lp[rows] - logp.gather(1, tgt[rows, None]).squeeze(1)).abs().max()) < 2e-3
    assert float((ent[rows] + (logp.exp() * logp).sum(-1)).abs().max()) < 2e-3
    # backward is linear in the coefficients; sum_j dlogits = -0 => dH of a pure-entropy-free uniform coef is finite
    c = torch.randn(M, device=d, generator=g) * 0.1
    zero = torch.zeros(M, device=d)
    gh1, gw1 = ops.lmhead_bwd(h, w, tgt, lse, ent, c, zero, 1.0)
    gh2, gw2 = ops.lmhead_bwd(h, w, tgt, lse, ent, 2 * c, zero, 1.0)
    assert _rel(gh2, 2 * gh1.float()) < 1e-2 and _rel(gw2, 2 * gw1.float()) < 1e-2
    # rows with a zero coefficient get exactly zero dH
    c0 = c.clone(); c0[::2] = 0
    gh3, _ = ops.lmhead_bwd(h, w, tgt, lse, ent, c0, zero, 1.0, need_grad_weight=False)
    assert torch.count_nonzero(gh3[::2]) == 0
    # dW check on a vocab slice against torch: dW[v] = sum_t dlogits[t,v] h[t]
    vs = slice(1000, 1256)
    p = torch.exp(h.float() @ w[vs].float().t() - lse[:, None])
    onehot = (tgt[:, None] == torch.arange(1000, 1256, device=d)[None]).float()
    dl = c[:, None] * (onehot - p)
    assert _rel(gw1[vs], dl.t() @ h.float()) < GRAD_RTOL


def test_rl_step_recompute_mode_matches_saved_mode():
    """set_save_logits(False): nothing of size tokens x vocab is kept between forward and backward."""
    from pipelinerl_swe_b200 import RLConfig, rl_step
    from pipelinerl_swe_b200.lm_head import set_save_logits
    H, V, lens = 128, 3000, [70, 41, 90]
    batch = synthetic_packed_batch(lens, V, seed=5)
    cfg = RLConfig(policy_loss="ppo", batch_size=3, kl_coef=0.1, final_kl_coef=0.1, entropy_bonus=0.01, final_entropy_bonus=0.01)
    grads = []
    try:
        for flag in (True, False):
            set_save_logits(flag)
            model = StubLM(H, V, sum(lens), seed=8).to(_dev())
            loss, stats = rl_step(model, _to_dev(batch), 0, 1, cfg)
            loss.backward()
            grads.append((loss.item(), model.lm_head.weight.grad.float(), model.hidden.grad.float()))
    finally:
        set_save_logits(True)
    assert grads[0][0] == grads[1][0]
    assert _rel(grads[0][1], grads[1][1]) < 1e-2 and _rel(grads[0][2], grads[1][2]) < 1e-2


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 1)])
def test_gemm_splitk_and_raster(a_mn, b_mn):
    from pipelinerl_swe_b200 import ops
    d = _dev()
    g = torch.Generator(device=d).manual_seed(1)
    M, N, K = (776 if a_mn else 777), 520, 9000  # MN-major operands need a 16-byte aligned row pitch (TMA)
    A = torch.randn(M, K, device=d, generator=g).bfloat16()
    B = torch.randn(N, K, device=d, generator=g).bfloat16()
    ref = A.float() @ B.float().t()
    Ain = A.t().contiguous() if a_mn else A
    Bin = B.t().contiguous() if b_mn else B
    for group_m, ks in ((0, 1), (1, 1), (2, 3), (3, 5)):
        out = ops.gemm_bf16(Ain, Bin, a_mn_major=bool(a_mn), b_mn_major=bool(b_mn), group_m=group_m, k_splits=ks)
        assert _rel(out, ref) < 1e-4, (group_m, ks)
        acc = torch.full((M, N), 2.0, device=d)
        ops.gemm_bf16(Ain, Bin, a_mn_major=bool(a_mn), b_mn_major=bool(b_mn), out=acc, accumulate=True, group_m=group_m, k_splits=ks)
        assert _rel(acc, ref + 2.0) < 1e-4


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_gemm_256_row_tiles_bit_equal_to_128_row_tiles(a_mn, b_mn):
    """gemm_big_m: two 128-row MMAs per CTA share one B tile; same k order per output => identical bits."""
    from pipelinerl_swe_b200 import _abi, ops
    d = _dev()
    g = torch.Generator(device=d).manual_seed(5)
    M, N, K = 1000, 776, 1192
    A = torch.randn(M, K, device=d, generator=g).bfloat16()
    B = torch.randn(N, K, device=d, generator=g).bfloat16()
    ref = A.float() @ B.float().t()
    Ain = A.t().contiguous() if a_mn else A
    Bin = B.t().contiguous() if b_mn else B
    try:
        outs = []
        for big in (0, 1):
            _abi.set_tunable("gemm_big_m", big)
            outs.append(ops.gemm_bf16(Ain, Bin, a_mn_major=bool(a_mn), b_mn_major=bool(b_mn), group_m=2, k_splits=1))
            outs.append(ops.gemm_bf16(Ain, Bin, a_mn_major=bool(a_mn), b_mn_major=bool(b_mn), group_m=2, k_splits=3))
    finally:
        _abi.set_tunable("gemm_big_m", 0)
    assert _rel(outs[2], ref) < 1e-4
    assert torch.equal(outs[0], outs[2]) and torch.equal(outs[1], outs[3])


@pytest.mark.parametrize("big", [0, 1])
@pytest.mark.parametrize("a_mn,b_mn,M,N,K", [(0, 0, 1000, 776, 1192),      # fewer tiles than clusters: every tile is cut
                                             (0, 1, 4000, 2560, 2048),     # whole waves + a cut tail
                                             (1, 1, 5000, 2056, 1536)])
def test_gemm_tail_split(a_mn, b_mn, M, N, K, big):
    """k_splits=0: tiles of the last partial wave are cut along K and summed in a fixed order (deterministic)."""
    from pipelinerl_swe_b200 import _abi, ops
    d = _dev()
    g = torch.Generator(device=d).manual_seed(6)
    A = torch.randn(M, K, device=d, generator=g).bfloat16()
    B = torch.randn(N, K, device=d, generator=g).bfloat16()
    ref = A.float() @ B.float().t()
    Ain = A.t().contiguous() if a_mn else A
    Bin = B.t().contiguous() if b_mn else B
    kw = dict(a_mn_major=bool(a_mn), b_mn_major=bool(b_mn), k_splits=0, group_m=3)
    try:
        _abi.set_tunable("gemm_big_m", big)
        f32 = ops.gemm_bf16(Ain, Bin, out_dtype=torch.float32, **kw)
        b16 = ops.gemm_bf16(Ain, Bin, out_dtype=torch.bfloat16, **kw)
        acc = torch.full((M, N), 2.0, device=d)
        ops.gemm_bf16(Ain, Bin, out=acc, accumulate=True, **kw)
        again = ops.gemm_bf16(Ain, Bin, out_dtype=torch.float32, **kw)
    finally:
        _abi.set_tunable("gemm_big_m", 0)
    assert _rel(f32, ref) < 1e-4 and _rel(b16, ref) < 1e-2 and _rel(acc, ref + 2.0) < 1e-4
    assert torch.equal(f32, again)


@pytest.mark.parametrize("dh_big,dw_big,tail", [(0, 0, 0), (1, 1, 1), (0, 1, 1), (1, 0, 0)])
def test_lmhead_bwd_tile_variants_agree(dh_big, dw_big, tail):
    """dH / dW through 128- or 256-row tiles, with or without the tail split, against torch autograd in fp32."""
    from pipelinerl_swe_b200 import _abi, ops
    d = _dev()
    M, H, V, temp = 1100, 768, 40000 + 8, 0.8
    g = torch.Generator(device=d).manual_seed(9)
    h = torch.randn(M, H, device=d, generator=g).bfloat16()
    w = (torch.randn(V, H, device=d, generator=g) * 0.04).bfloat16()
    tgt = torch.randint(0, V, (M,), device=d, generator=g)
    c_lp = torch.randn(M, device=d, generator=g) * 0.1
    c_ent = torch.randn(M, device=d, generator=g) * 0.01
    hr, wr = h.float().requires_grad_(True), w.float().requires_grad_(True)
    logp = torch.log_softmax((hr @ wr.t()) / temp, -1)
    lp = logp.gather(1, tgt[:, None]).squeeze(1)
    ent = -(logp.exp() * logp).sum(-1)
    ((lp * c_lp).sum() + (ent * c_ent).sum()).backward()
    try:
        _abi.set_tunable("dh_big_m", dh_big)
        _abi.set_tunable("dw_big_m", dw_big)
        _abi.set_tunable("tail_split", tail)
        saved = ops.lmhead_saved_buffer(M, H, V, d)
        stats = ops.lmhead_fwd_stats(h, w, tgt, 1.0 / temp, saved=saved)
        lp2, ent2, lse = ops.lmhead_finalize(stats)
        gw = torch.full((V, H), 0.5, device=d)
        gh, _ = ops.lmhead_bwd(h, w, tgt, lse, ent2, c_lp, c_ent, 1.0 / temp, grad_weight_out=gw, accumulate=True, saved=saved)
    finally:
        _abi.set_tunable("dh_big_m", 1)
        _abi.set_tunable("dw_big_m", 0)
        _abi.set_tunable("tail_split", 1)
    assert _rel(gh, hr.grad) < GRAD_RTOL
    assert _rel(gw - 0.5, wr.grad) < GRAD_RTOL


@pytest.mark.parametrize("entropy", [True, False])
def test_soft_k_lockstep_never_changes_a_bit(entropy):
    """Pacing of the TMA producers (lockstep_fwd / _dh / _dw, any epoch length or slack, or off) is scheduling only: forward
    statistics, dH and the accumulated f32 dW are bit-identical whatever the setting.  Shape with several waves, a long K for dH
    and the tail split."""
    from pipelinerl_swe_b200 import _abi, ops
    d = _dev()
    M, H, V, temp = 2100, 1024, 60000 + 8, 0.9
    g = torch.Generator(device=d).manual_seed(19)
    h = torch.randn(M, H, device=d, generator=g).bfloat16()
    w = (torch.randn(V, H, device=d, generator=g) * 0.04).bfloat16()
    tgt = torch.randint(0, V, (M,), device=d, generator=g)
    c_lp = torch.randn(M, device=d, generator=g) * 0.1
    c_ent = torch.randn(M, device=d, generator=g) * 0.01 if entropy else None
    names = ("lockstep_fwd", "lockstep_dh", "lockstep_dw", "lockstep_slack")
    shipped = [_abi.set_tunable(n, 0) for n in names]          # set_tunable returns the previous value
    outs = []
    try:
        for setting in ((0, 0, 0, 2), tuple(shipped), (4, 4, 4, 1), (2, 16, 8, 3), (1, 1, 1, 1)):
            for n, v in zip(names, setting):
                _abi.set_tunable(n, v)
            saved = ops.lmhead_saved_buffer(M, H, V, d)
            stats = ops.lmhead_fwd_stats(h, w, tgt, 1.0 / temp, saved=saved)
            lp, ent, lse = ops.lmhead_finalize(stats)
            gw = torch.full((V, H), 0.25, device=d)
            gh, _ = ops.lmhead_bwd(h, w, tgt, lse, ent, c_lp, c_ent, 1.0 / temp, grad_weight_out=gw, accumulate=True, saved=saved)
            gh2, gw2 = ops.lmhead_bwd(h, w, tgt, lse, ent, c_lp, c_ent, 1.0 / temp)      # recompute mode, bf16 dW
            torch.cuda.synchronize()
            outs.append((stats.clone(), gh.clone(), gw, gh2.clone(), gw2.clone()))
    finally:
        for n, v in zip(names, shipped):
            _abi.set_tunable(n, v)
    assert shipped[0] > 0 and shipped[1] > 0     # the shipped schedule paces
    for o in outs[1:]:
        for a, b in zip(o, outs[0]):
            assert torch.equal(a, b)


def test_lmhead_forward_backward_replays_from_a_cuda_graph():
    """K1 -> finalize -> K2 captured into a CUDA graph (after one warm-up call on the capture stream) and replayed: bit-equal to
    the eager call every time.  The library allocates nothing, its per-launch lockstep counters are a memset node of the graph."""
    from pipelinerl_swe_b200 import ops
    d = _dev()
    M, H, V = 1100, 768, 40000 + 8
    g = torch.Generator(device=d).manual_seed(3)
    h = torch.randn(M, H, device=d, generator=g).bfloat16()
    w = (torch.randn(V, H, device=d, generator=g) * 0.04).bfloat16()
    tgt = torch.randint(0, V, (M,), device=d, generator=g)
    c_lp = torch.randn(M, device=d, generator=g) * 0.1

    def step():
        saved = ops.lmhead_saved_buffer(M, H, V, d)
        stats = ops.lmhead_fwd_stats(h, w, tgt, 1.0, saved=saved)
        lp, ent, lse = ops.lmhead_finalize(stats)
        gh, gw = ops.lmhead_bwd(h, w, tgt, lse, ent, c_lp, None, 1.0, saved=saved)
        return lp, ent, gh, gw

    eager = [t.clone() for t in step()]
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step()
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out = step()
    for _ in range(3):
        for t in out:
            t.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert all(torch.equal(a, b) for a, b in zip(out, eager))


@pytest.mark.parametrize("f32_out", [True, False])
@pytest.mark.parametrize("scale", [8.0, 12.0, 16.0])
def test_entropy_free_backward_on_confident_tokens(scale, f32_out):
    """Rows whose target is the peak of the distribution (median p_target 0.34 / 0.96 / 0.999): the gradient a (onehot - p) is a
    difference of nearly equal numbers at the target.  The saved-exponentials backward folds the onehot into the target's
    exponential in f32 BEFORE the GEMMs, so dH and dW keep the north-star tolerance against fp32 autograd, bf16 or f32 output
    (adding the onehot terms after a bf16-rounded GEMM result was 7 % / 140 % off at p = 0.96 / 0.999)."""
    from pipelinerl_swe_b200 import ops
    d = _dev()
    M, H, V = 512, 256, 5000
    g = torch.Generator(device=d).manual_seed(5)
    w = torch.randn(V, H, device=d, generator=g)
    w = (w / w.norm(dim=1, keepdim=True)).bfloat16()
    tgt = torch.randint(0, V, (M,), device=d, generator=g)
    c_lp = torch.randn(M, device=d, generator=g) * 0.1
    h = (w[tgt].float() * scale + 0.3 * torch.randn(M, H, device=d, generator=g)).bfloat16()
    hr, wr = h.float().requires_grad_(True), w.float().requires_grad_(True)
    lp = torch.log_softmax(hr @ wr.t(), -1).gather(1, tgt[:, None]).squeeze(1)
    (lp * c_lp).sum().backward()
    for save in (True, False):
        saved = ops.lmhead_saved_buffer(M, H, V, d) if save else None
        stats = ops.lmhead_fwd_stats(h, w, tgt, 1.0, saved=saved)
        lp2, ent2, lse = ops.lmhead_finalize(stats)
        gw0 = torch.zeros(V, H, device=d) if f32_out else None
        gh, gw = ops.lmhead_bwd(h, w, tgt, lse, ent2, c_lp, None if save else torch.zeros_like(c_lp), 1.0, saved=saved,
                                grad_weight_out=gw0, accumulate=bool(f32_out))
        assert _rel(lp2, lp.detach()) < LP_RTOL
        assert _rel(gh, hr.grad) < GRAD_RTOL, (save, _rel(gh, hr.grad))
        assert _rel(gw, wr.grad) < GRAD_RTOL, (save, _rel(gw, wr.grad))


class _EmbedLM(torch.nn.Module):
    """Token-wise 'trunk' (embedding lookup) + lm_head: hidden depends on the ids, so packing / offsets are exercised."""

    def __init__(self, H, V, seed=0):
        super().__init__()
        g = torch.Generator().manual_seed(seed)
        self.embed = torch.nn.Embedding(V, H, dtype=torch.bfloat16)
        self.lm_head = torch.nn.Linear(H, V, bias=False, dtype=torch.bfloat16)
        with torch.no_grad():
            self.embed.weight.copy_(torch.randn(V, H, generator=g).bfloat16())
            self.lm_head.weight.copy_((torch.randn(V, H, generator=g) * 0.05).bfloat16())
        self.seen_rows = []

    def forward(self, input_ids=None, attention_mask=None, position_ids=None, **kw):
        import types
        self.seen_rows.append((input_ids.shape[1], position_ids[0].tolist()))
        return types.SimpleNamespace(logits=self.lm_head(self.embed(input_ids)))


def test_ref_logprob_annotation_matches_fp32_log_softmax():
    """preprocess.py:73-91 semantics: ref_logprobs[j] = log p(completion token j | everything before it), T = 1."""
    from pipelinerl_swe_b200.ref_logprobs import annotate_traces_with_ref_logprobs
    H, V = 256, 5000 + 8
    model = _EmbedLM(H, V).to(_dev())
    g = torch.Generator().manual_seed(4)
    traces = []
    for n, c in [(30, 10), (200, 199), (64, 1), (150, 70), (90, 33), (400, 20)]:
        traces.append({"input_ids": torch.randint(0, V, (n,), generator=g).tolist(), "logprobs": [0.0] * c})
    annotate_traces_with_ref_logprobs(model, traces, max_tokens=300)
    assert [r[0] for r in model.seen_rows] == [294, 240, 400]          # 30+200+64 | 150+90 | 400 (alone, over the cap)
    assert model.seen_rows[0][1][:31] == list(range(30)) + [0]          # position ids restart per trace
    W, E = model.lm_head.weight.detach().float().cpu(), model.embed.weight.detach().float().cpu()
    for tr in traces:
        ids = torch.tensor(tr["input_ids"])
        logp = torch.log_softmax(E[ids] @ W.t(), -1)
        c = len(tr["logprobs"])
        want = logp[len(ids) - c - 1: len(ids) - 1].gather(1, ids[len(ids) - c:, None]).squeeze(1)
        got = torch.tensor(tr["ref_logprobs"])
        assert got.shape == want.shape
        assert float((got - want).abs().max()) <= LP_RTOL * float(want.abs().max())
    assert isinstance(model.lm_head, torch.nn.Module) and model.training   # mode restored
    with pytest.raises(ValueError):
        annotate_traces_with_ref_logprobs(model, [{"input_ids": [1, 2], "logprobs": [0.0, 0.0]}])


@pytest.mark.parametrize("seed", [100, 101])
def test_gemm_random_shapes_layouts_and_split_modes(seed):
    """24 random (M, N, K, layout, 1/2-CTA, 128/256-row tiles, uniform / tail split, accumulate) cases vs fp32 matmul."""
    from gpu_probe import probe_gemm_fuzz
    ok, msg = probe_gemm_fuzz(seed)
    assert ok, msg


@pytest.mark.parametrize("seed", [200, 201])
def test_lmhead_random_shapes_and_variants(seed):
    """16 random (M, H, V, temperature, tile variant, saved / recompute) forward + backward cases vs torch autograd fp32."""
    from gpu_probe import probe_lmhead_fuzz
    ok, msg = probe_lmhead_fuzz(seed)
    assert ok, msg


@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_rl_step_with_master_weights_in_another_dtype(dtype, fused):
    """fp32 (torch AMP / accelerate mixed precision) or fp16 parameters: the GEMMs read a bf16 cast, exactly what a bf16 model
    holding those cast values computes; gradients come back in the parameter's dtype, or -- fused accumulation -- as f32
    ``main_grad`` on the master weight with no bf16 [V,H] temporary."""
    from pipelinerl_swe_b200 import RLConfig, fold_main_grad, rl_step, set_fused_grad_accumulation
    T, H, V = 300, 136, 1032
    batch = _to_dev(synthetic_packed_batch([100, 120, 80], V, seed=11))
    cfg = RLConfig(batch_size=4, kl_coef=0.05, final_kl_coef=0.05)
    md = StubLM(H, V, T, seed=3, dtype=dtype).to(_dev())
    mb = StubLM(H, V, T, seed=3).to(_dev())
    with torch.no_grad():
        mb.hidden.copy_(md.hidden.to(torch.bfloat16))
        mb.lm_head.weight.copy_(md.lm_head.weight.to(torch.bfloat16))
    set_fused_grad_accumulation(fused)
    try:
        out = []
        for m in (md, mb):
            for _ in range(2):                       # two micro-batches accumulate
                loss, stats = rl_step(m, batch, 0, 1, cfg)
                loss.backward()
            out.append((loss.item(), stats))
        assert out[0][0] == out[1][0] and out[0][1] == out[1][1]      # same kernels on the same bf16 values
        assert md.hidden.grad.dtype == dtype
        assert _rel(md.hidden.grad, mb.hidden.grad) < 2e-3            # the bf16 dH cast to the parameter's dtype
        if fused:
            assert md.lm_head.weight.grad is None and md.lm_head.weight.main_grad.dtype == torch.float32
            assert torch.equal(md.lm_head.weight.main_grad, mb.lm_head.weight.main_grad)
            fold_main_grad(md)
            assert md.lm_head.weight.grad.dtype == dtype and float(md.lm_head.weight.main_grad.abs().max()) == 0.0
            assert _rel(md.lm_head.weight.grad, mb.lm_head.weight.main_grad) < (1e-6 if dtype == torch.float32 else 2e-3)
        else:
            assert md.lm_head.weight.grad.dtype == dtype
            assert _rel(md.lm_head.weight.grad, mb.lm_head.weight.grad) < 1e-2   # two bf16 micro-batch gradients summed in dtype vs in bf16
    finally:
        set_fused_grad_accumulation(False)


@pytest.mark.parametrize("seed", [300, 301, 302])
def test_rl_step_random_configs_vs_the_reference_rl_step(seed):
    """Whole drop-in (rl_step + backward) against the reference's own rl_step (oracle/_ref) over random batches and RLConfigs."""
    from fuzz_rl_step import fuzz
    ok, msg = fuzz(seed, n_cases=10)
    print(msg)
    assert ok, msg


def test_rl_step_autograd_usage_patterns_match_the_reference():
    """Scaled loss, two graphs alive + one backward, backward twice, no_grad evaluation, autograd.grad w.r.t. hidden only, a side
    stream: same gradients as the reference's rl_step under the same use."""
    from fuzz_rl_step import usage_patterns
    res = usage_patterns()
    print(res)
    assert all(ok for _, ok, _ in res), [r for r in res if not r[1]]


def test_rl_step_raises_what_the_reference_raises():
    """Non-finite inputs, group_tokens <= 0, unknown policy_loss: the same exception TYPE as the reference's rl_step (rl/__init__.py:209-313)."""
    from fuzz_rl_step import check_error_case, error_cases
    for name, mutate in error_cases():
        ref_exc, our_exc = check_error_case(name, mutate)
        assert ref_exc is not None, f"{name}: the reference did not raise"
        assert our_exc is not None and issubclass(our_exc, ref_exc), f"{name}: reference raised {ref_exc}, drop-in raised {our_exc}"


@pytest.mark.parametrize("tied", [False, True])
def test_rl_step_through_a_real_hf_causal_lm(tied):
    """Drop-in check on an unmodified transformers model (random tiny Qwen2): the lm_head swap inside HF's own forward,
    autograd from the fused op back into the trunk, and tied embeddings, against the reference math on the plain logits."""
    transformers = pytest.importorskip("transformers")
    from pipelinerl_swe_b200 import FusedLMHead, RLConfig, rl_step
    d = _dev()
    V, lens = 2048 + 8, [60, 41, 100]
    T = sum(lens)
    cfg_m = transformers.Qwen2Config(vocab_size=V, hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=4,
                                     num_key_value_heads=2, max_position_embeddings=512, tie_word_embeddings=tied)
    torch.manual_seed(3)
    model = transformers.Qwen2ForCausalLM(cfg_m).to(torch.bfloat16).to(d)
    batch = synthetic_packed_batch(lens, V, seed=5, n_prompt=6)
    inputs = dict(input_ids=batch["input_ids"].to(d), attention_mask=batch["attention_mask"].to(d), position_ids=batch["position_ids"].to(d))
    cfgkw = dict(policy_loss="ppo", batch_size=3, kl_coef=0.1, final_kl_coef=0.1, temperature=0.8, epsilon=0.2,
                 entropy_bonus=0.01, final_entropy_bonus=0.01)

    # reference math on the plain HF logits (what rl/__init__.py:197-208 does), old/ref log-probs near the policy
    out = model(**inputs, output_hidden_states=True)
    logits = out.logits.float()[0, :-1] / cfgkw["temperature"]
    logp = torch.log_softmax(logits, -1)
    new_lp = logp.gather(1, inputs["input_ids"][0, 1:, None]).squeeze(1)
    ent = -(logp.exp() * logp).sum(-1)
    g = torch.Generator().manual_seed(1)
    m = (batch["labels"][:, 1:] != -100).float()
    batch["old_logprobs"][:, 1:] = (new_lp.detach().cpu()[None] + 0.05 * torch.randn(1, T - 1, generator=g)) * m
    batch["ref_logprobs"][:, 1:] = (new_lp.detach().cpu()[None] + 0.10 * torch.randn(1, T - 1, generator=g)) * m
    bd = _to_dev(batch)
    want_loss, want_stats = oracle.oracle_token_loss(new_lp[None], ent[None], bd, 1, 10, oracle.OracleRLConfig(**cfgkw))
    want_loss.backward()
    ref_grads = {n: p.grad.detach().float().clone() for n, p in model.named_parameters() if p.grad is not None}
    model.zero_grad(set_to_none=True)

    loss, stats = rl_step(model, bd, 1, 10, RLConfig(**cfgkw))
    assert isinstance(model.lm_head, FusedLMHead)
    if tied:
        assert model.lm_head.weight is model.model.embed_tokens.weight
    loss.backward()
    assert abs(loss.item() - want_loss.item()) <= 5e-3 * abs(want_loss.item()) + 1e-5   # bf16 logits on the reference side
    assert_stats_close(stats, want_stats, rtol=1e-2, atol=2e-3)
    for name in ("lm_head.weight", "model.norm.weight", "model.layers.1.mlp.down_proj.weight", "model.layers.0.self_attn.q_proj.weight",
                 "model.embed_tokens.weight"):
        if name not in ref_grads:
            continue
        got = dict(model.named_parameters())[name].grad.float()
        assert _rel(got, ref_grads[name]) < 5e-2, name   # both sides carry bf16 rounding through the trunk
    # outside rl_step the swapped head is a plain projection again
    plain = model(**inputs).logits
    assert plain.shape == (1, T, V)


def test_rl_step_through_the_value_head_wrapper():
    """The reference's AutoModelForCausalLMWithValueHead shape (finetune/value_model.py:55-116): lm_head lives at
    ``.pretrained_model.lm_head``, values come from ``Linear(H, 1)`` on the last hidden state."""
    transformers = pytest.importorskip("transformers")
    import types
    from pipelinerl_swe_b200 import FusedLMHead, RLConfig, rl_step
    d = _dev()
    V, lens = 1024 + 8, [50, 33, 70]
    cfg_m = transformers.Qwen2Config(vocab_size=V, hidden_size=64, intermediate_size=128, num_hidden_layers=1, num_attention_heads=4,
                                     num_key_value_heads=2, max_position_embeddings=256, tie_word_embeddings=False)

    class WithValueHead(torch.nn.Module):
        def __init__(self, lm):
            super().__init__()
            self.pretrained_model = lm
            self.value_head = torch.nn.Linear(lm.config.hidden_size, 1)

        def forward(self, **kw):
            out = self.pretrained_model(**kw, output_hidden_states=True)
            value = self.value_head(out.hidden_states[-1].float()).squeeze(-1)
            return types.SimpleNamespace(logits=out.logits, value=value)

    torch.manual_seed(4)
    model = WithValueHead(transformers.Qwen2ForCausalLM(cfg_m).to(torch.bfloat16)).to(d)
    batch = _to_dev(synthetic_packed_batch(lens, V, seed=6, n_prompt=5))
    cfg = RLConfig(batch_size=3, kl_coef=0.0, final_kl_coef=0.0, value_loss_coef=0.5, clamp_log_ratio_ref_new_value=5)
    loss, stats = rl_step(model, batch, 0, 1, cfg)
    assert isinstance(model.pretrained_model.lm_head, FusedLMHead)
    assert {"value_mean", "value_loss", "value_mse", "value_max", "value_min"} <= set(stats)
    loss.backward()
    assert model.value_head.weight.grad is not None and float(model.value_head.weight.grad.abs().sum()) > 0
    assert model.pretrained_model.lm_head.weight.grad is not None
    assert torch.isfinite(loss).item()


@pytest.mark.parametrize("save", [False, True])
def test_lmhead_bwd_in_two_stages_is_bit_identical(save):
    """stages = DLOGITS|DH then DW (what the vocab-parallel path does around its all-reduce) == the single call."""
    from pipelinerl_swe_b200 import _abi, ops
    d = _dev()
    M, H, V, temp = 700, 256, 30000 + 8, 0.9
    g = torch.Generator(device=d).manual_seed(12)
    h = torch.randn(M, H, device=d, generator=g).bfloat16()
    w = (torch.randn(V, H, device=d, generator=g) * 0.05).bfloat16()
    tgt = torch.randint(0, V, (M,), device=d, generator=g)
    c = torch.randn(2, M, device=d, generator=g) * 0.1
    outs = []
    for staged in (False, True):
        saved = ops.lmhead_saved_buffer(M, H, V, d) if save else None
        lp, ent, lse = ops.lmhead_finalize(ops.lmhead_fwd_stats(h, w, tgt, 1.0 / temp, saved=saved))
        args = (h, w, tgt, lse, ent, c[0].contiguous(), c[1].contiguous(), 1.0 / temp)
        if staged:
            gh, none_w = ops.lmhead_bwd(*args, saved=saved, stages=_abi.BWD_DLOGITS | _abi.BWD_DH)
            none_h, gw = ops.lmhead_bwd(*args, saved=saved, stages=_abi.BWD_DW)
            assert none_w is None and none_h is None
        else:
            gh, gw = ops.lmhead_bwd(*args, saved=saved)
        outs.append((gh.clone(), gw.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.gpu
@pytest.mark.parametrize("with_value_head", [False, True])
def test_fused_final_norm_inside_an_unmodified_transformers_model(with_value_head):
    """install_fused_final_norm on a random tiny Qwen2 (and the reference's value-head wrapper shape, value_model.py:40-116):
    same loss / stats / gradients as the model's own RMSNorm + Linear, module identity and state dict untouched."""
    transformers = pytest.importorskip("transformers")
    import copy
    import types
    from pipelinerl_swe_b200 import RLConfig, install_fused_final_norm, rl_step
    from pipelinerl_swe_b200.final_norm import FusedFinalNorm, FusedValueHead
    d = _dev()
    V, lens = 1024 + 8, [50, 33, 70]
    cfg_m = transformers.Qwen2Config(vocab_size=V, hidden_size=64, intermediate_size=128, num_hidden_layers=1, num_attention_heads=4,
                                     num_key_value_heads=2, max_position_embeddings=256, tie_word_embeddings=False)

    class ValueHead(torch.nn.Module):          # value_model.py:40-52
        def __init__(self, H):
            super().__init__()
            self.output = torch.nn.Linear(H, 1)

        def forward(self, hidden_states):
            return self.output(hidden_states).squeeze(-1)

    class WithValueHead(torch.nn.Module):      # value_model.py:55-116
        def __init__(self, lm):
            super().__init__()
            self.pretrained_model = lm
            self.value_head = ValueHead(lm.config.hidden_size)

        def forward(self, **kw):
            out = self.pretrained_model(**kw, output_hidden_states=True)
            return types.SimpleNamespace(logits=out.logits, value=self.value_head(out.hidden_states[-1]))

    torch.manual_seed(4)
    lm = transformers.Qwen2ForCausalLM(cfg_m)
    with torch.no_grad():
        lm.model.norm.weight.add_(0.2 * torch.randn_like(lm.model.norm.weight))
    plain = (WithValueHead(lm) if with_value_head else lm).to(torch.bfloat16).to(d)
    fused = copy.deepcopy(plain)
    keys_before = list(fused.state_dict().keys())
    norm = install_fused_final_norm(fused)
    assert isinstance(norm, FusedFinalNorm) and install_fused_final_norm(fused) is norm        # idempotent
    assert list(fused.state_dict().keys()) == keys_before
    if with_value_head:
        assert isinstance(fused.value_head, FusedValueHead)
    batch = _to_dev(synthetic_packed_batch(lens, V, seed=6, n_prompt=5))
    cfg = RLConfig(batch_size=3, kl_coef=0.05, final_kl_coef=0.05, value_loss_coef=0.5, clamp_log_ratio_ref_new_value=5)
    out = []
    for m in (plain, fused):
        loss, stats = rl_step(m, batch, 0, 1, cfg)
        loss.backward()
        out.append((loss.item(), stats, {n: p.grad.float().clone() for n, p in m.named_parameters() if p.grad is not None}))
    (l0, s0, g0), (l1, s1, g1) = out
    assert abs(l0 - l1) <= 1e-3 * abs(l0) + 1e-5, (l0, l1)
    assert_stats_close(s1, s0, rtol=5e-3, atol=5e-4)
    assert set(g0) == set(g1)
    for n in g0:
        err = (g0[n] - g1[n]).abs().max().item() / max(g0[n].abs().max().item(), 1e-12)
        assert err < 3e-2, (n, err)   # bf16 autograd on both sides


def test_fused_grad_accumulation_three_micro_batches_sum_in_f32():
    """set_fused_grad_accumulation(True): the dW GEMM epilogue adds into lm_head.weight.main_grad (f32); three micro-batches
    accumulate to the f32 sum of their gradients, autograd sees no weight gradient, fold_main_grad hands it to .grad."""
    from pipelinerl_swe_b200 import RLConfig, fold_main_grad, rl_step, set_fused_grad_accumulation
    H, V = 128, 3000
    cfg = RLConfig(policy_loss="ppo", batch_size=3, kl_coef=0.1, final_kl_coef=0.1, entropy_bonus=0.01, final_entropy_bonus=0.01)
    cases = [([70, 41, 90], 5), ([201], 6), ([30, 30, 60, 81], 7)]
    # per-micro-batch gradients through the default path (fresh dW each), summed in f32 on the host side
    want = torch.zeros(V, H, device=_dev())
    models = []
    for lens, seed in cases:
        m = StubLM(H, V, sum(lens), seed=8).to(_dev())          # same lm_head weights (seed), different hidden length
        loss, _ = rl_step(m, _to_dev(synthetic_packed_batch(lens, V, seed=seed)), 0, 1, cfg)
        loss.backward()
        want += m.lm_head.weight.grad.float()
        models.append(m)
    set_fused_grad_accumulation(True)
    try:
        head_owner = StubLM(H, V, 201, seed=8).to(_dev())
        for (lens, seed), m in zip(cases, models):
            m.lm_head = head_owner.lm_head                       # ONE shared lm_head parameter across the micro-batches
            m.zero_grad(set_to_none=True)
            loss, _ = rl_step(m, _to_dev(synthetic_packed_batch(lens, V, seed=seed)), 0, 1, cfg)
            loss.backward()
            assert m.hidden.grad is not None
        w = head_owner.lm_head.weight
        assert w.grad is None and w.main_grad.dtype == torch.float32
        assert _rel(w.main_grad, want) < 5e-3                     # `want` carries one bf16 rounding per micro-batch
        fold_main_grad(head_owner)
        assert w.grad is not None and w.grad.dtype == torch.bfloat16 and _rel(w.grad, want) < 1e-2
        assert torch.count_nonzero(w.main_grad) == 0
    finally:
        set_fused_grad_accumulation(False)


@pytest.mark.parametrize("save", [True, False])
def test_row_chunked_backward_under_a_scratch_budget(save):
    """A 20 000-token row at the real vocab size exceeds a 2 GiB [rows x vocab] budget: forward + backward run in row chunks
    (first chunk keeps its logits, the rest are recomputed) and reproduce the un-chunked result; the scratch stays under
    the budget."""
    from pipelinerl_swe_b200 import lm_head as LH
    from pipelinerl_swe_b200 import ops
    d = _dev()
    M, H, V, temp = 20000, 256, 152064, 0.9
    g = torch.Generator(device=d).manual_seed(21)
    h = torch.randn(M, H, device=d, generator=g).bfloat16().requires_grad_(True)
    w = (torch.randn(V, H, device=d, generator=g) * 0.05).bfloat16().requires_grad_(True)
    tgt = torch.randint(0, V, (M,), device=d, generator=g)
    c = torch.randn(M, 2, device=d, generator=g) * torch.tensor([0.1, 0.01], device=d)
    outs = []
    old_budget, old_save = LH.saved_logits_budget(), LH.save_logits_enabled()
    try:
        LH.set_save_logits(save)
        for budget in (64 << 30, 2 << 30):
            LH.set_saved_logits_budget(budget)
            n_chunks = len(LH.row_chunks(M, V))
            assert (n_chunks == 1) if budget > (8 << 30) else (n_chunks == 4)   # 2 GiB / 304128 B per row -> 6656-row chunks
            ops._ws_cache.clear()
            torch.cuda.empty_cache()
            torch.cuda.reset_peak_memory_stats(d)
            base = torch.cuda.memory_allocated(d)
            out = LH.LMHeadLogprobFunction.apply(h, w, tgt, 1.0 / temp, None, 0)
            (out * c).sum().backward()
            peak = torch.cuda.max_memory_allocated(d) - base
            outs.append((out.detach().clone(), h.grad.clone(), w.grad.clone(), peak))
            h.grad = w.grad = None
    finally:
        LH.set_saved_logits_budget(old_budget)
        LH.set_save_logits(old_save)
    (o0, gh0, gw0, p0), (o1, gh1, gw1, p1) = outs
    assert torch.equal(o0, o1)                                   # K1 is row-independent: bit-identical log-probs / entropies
    assert _rel(gh1, gh0.float()) < 1e-2 and _rel(gw1, gw0.float()) < 1e-2
    per_row = -(-V // 256) * 512
    assert p0 > M * per_row                                      # un-chunked: the whole [rows x vocab] bf16 buffer existed
    assert p1 < (2 << 30) * (2 if save else 1) + 3 * V * H * 4 + (64 << 20), (p0, p1)   # chunked: <= budget per live buffer + dW temporaries


def test_sp_slices_with_carried_boundary_targets_cover_the_unsliced_row():
    """Opt-in beyond the reference: make_slices(n, carry_boundary_target=True).  PPO with 1/batch_size token weights is a plain
    sum over target tokens, so the per-slice losses / token counts / gradients must add up to the un-sliced row's; without
    the carry (the reference's behaviour, default) the n-1 boundary targets are missing."""
    from pipelinerl_swe_b200 import PipelineBatchEncoding, RLConfig, rl_step
    H, V, lens, n_slices = 128, 3000, [70, 41, 90, 55], 4
    T = sum(lens)
    batch = synthetic_packed_batch(lens, V, seed=5, n_prompt=3)
    cfg = RLConfig(policy_loss="ppo", batch_size=4, kl_coef=0.1, final_kl_coef=0.1, entropy_bonus=0.01, final_entropy_bonus=0.01)
    full = StubLM(H, V, T, seed=8).to(_dev())
    loss_full, stats_full = rl_step(full, _to_dev(batch), 0, 1, cfg)
    loss_full.backward()
    pbe = PipelineBatchEncoding(**batch, model_version=0)
    results = {}
    for carry in (False, True):
        tot_loss, tot_tokens, gw = 0.0, 0, torch.zeros(V, H, device=_dev())
        gh = []
        for k, sl in enumerate(pbe.make_slices(n_slices, carry_boundary_target=carry)):
            a = k * T // n_slices
            m = StubLM(H, V, T // n_slices, seed=8).to(_dev())
            with torch.no_grad():
                m.hidden.copy_(full.hidden[:, a:a + T // n_slices])
                m.lm_head.weight.copy_(full.lm_head.weight)      # StubLM draws the head after the hidden states: same seed, other T
            loss, stats = rl_step(m, sl.to_device(_dev()), 0, 1, cfg)
            loss.backward()
            tot_loss += loss.item()
            tot_tokens += stats["num_output_tokens_sum"]
            gw += m.lm_head.weight.grad.float()
            gh.append(m.hidden.grad.float())
        results[carry] = (tot_loss, tot_tokens, gw, torch.cat(gh, dim=1))
    n_boundary_labelled = sum(int(batch["labels"][0, k * T // n_slices] != -100) for k in range(1, n_slices))
    assert n_boundary_labelled > 0
    assert results[False][1] == stats_full["num_output_tokens_sum"] - n_boundary_labelled      # the reference drops them
    assert results[True][1] == stats_full["num_output_tokens_sum"]                              # the carry restores them
    assert abs(results[True][0] - loss_full.item()) <= 1e-4 * abs(loss_full.item()) + 1e-6, (results[True][0], results[False][0], loss_full.item())
    assert abs(results[False][0] - loss_full.item()) > abs(results[True][0] - loss_full.item())
    assert _rel(results[True][2], full.lm_head.weight.grad.float()) < 1e-2
    assert _rel(results[True][3], full.hidden.grad.float()) < 1e-2


@pytest.mark.parametrize("f32", [True, False])
def test_entropy_free_backward_is_deterministic_with_repeated_targets(f32):
    """Hot tokens (many rows share a target) through the sparse target-row kernel: two runs give identical bits, in the f32
    accumulate form and in the bf16 form, and match the fp32 reference."""
    from pipelinerl_swe_b200 import ops
    d = _dev()
    M, H, V, temp = 2500, 520, 6000 + 8, 0.8
    g = torch.Generator(device=d).manual_seed(31)
    h = torch.randn(M, H, device=d, generator=g).bfloat16()
    w = (torch.randn(V, H, device=d, generator=g) * 0.05).bfloat16()
    tgt = torch.randint(0, 7, (M,), device=d, generator=g) * 811          # seven distinct targets, ~350 rows each
    tgt[::5] = torch.randint(0, V, (M,), device=d, generator=g)[::5]      # plus singletons
    c_lp = torch.randn(M, device=d, generator=g) * 0.1
    c_lp[::4] = 0.0
    hr, wr = h.float().requires_grad_(True), w.float().requires_grad_(True)
    logp = torch.log_softmax((hr @ wr.t()) / temp, -1)
    (logp.gather(1, tgt[:, None]).squeeze(1) * c_lp).sum().backward()
    outs = []
    for _ in range(2):
        saved = ops.lmhead_saved_buffer(M, H, V, d)
        lp, ent, lse = ops.lmhead_finalize(ops.lmhead_fwd_stats(h, w, tgt, 1.0 / temp, saved=saved))
        gw0 = torch.full((V, H), 0.25, device=d) if f32 else None
        gh, gw = ops.lmhead_bwd(h, w, tgt, lse, ent, c_lp, None, 1.0 / temp, saved=saved, grad_weight_out=gw0, accumulate=f32)
        outs.append((gh.clone(), gw.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    got = outs[0][1] - 0.25 if f32 else outs[0][1]
    assert _rel(got, wr.grad) < (1e-3 if f32 else 1e-2) and _rel(outs[0][0], hr.grad) < GRAD_RTOL


def test_forward_reference_range_is_exact_to_60_nats_and_fails_loudly_beyond_88():
    """K1 takes its exponentials relative to a per-row reference r = max(target logit, four sampled logits).  A target 60 or
    even 100 nats below the row maximum is exact as long as ONE of the five logits is within 88 nats of the maximum (here the
    sampled columns are, at logit 0).  Only when the target AND every sampled column sit more than 88 nats below the maximum
    (probabilities below 1e-38, outside f32's exponent range in these units) does the row give a NON-FINITE log-prob -- which
    rl_step turns into the reference's own assert (rl/__init__.py:209) -- never a silently wrong number.  Neighbouring rows are
    unaffected in every case."""
    from pipelinerl_swe_b200 import ops
    d = _dev()
    M, H, V = 64, 64, 4096
    g = torch.Generator(device=d).manual_seed(3)
    u = torch.zeros(H, device=d)
    u[0] = 1.0
    w = (torch.randn(V, H, device=d, generator=g) * 0.05)
    w[:, 0] = 0.0
    hot, tgt_tok, sampled = 7, 11, [0, V // 4, V // 2, 3 * V // 4]
    for gap_half, sink_samples, finite in ((30.0, False, True), (50.0, False, True), (50.0, True, False)):
        ww = w.clone()
        ww[hot] = gap_half * u                                # logit +gap/2 on the hot token ...
        ww[tgt_tok] = -gap_half * u                           # ... and -gap/2 on the target of row 0
        if sink_samples:
            ww[sampled] = -gap_half * u                       # contrived: the sampled columns are as unlikely as the target
        h = torch.randn(M, H, device=d, generator=g) * 0.5
        h[0] = u                                              # row 0 sees the extreme pair; the other rows see a zero first component
        h[1:, 0] = 0.0
        tgt = torch.randint(0, V, (M,), device=d, generator=g)
        tgt[0] = tgt_tok
        hb, wb = h.bfloat16(), ww.bfloat16()
        lp, ent, lse = ops.lmhead_finalize(ops.lmhead_fwd_stats(hb, wb, tgt, 1.0))
        ref = torch.log_softmax(hb.float() @ wb.float().t(), -1).gather(1, tgt[:, None]).squeeze(1)
        assert float((lp[1:] - ref[1:]).abs().max()) < 1e-3   # neighbours are exact either way
        if finite:
            assert abs(float(lp[0] - ref[0])) < 1e-3 * abs(float(ref[0])) and float(ref[0]) < -55.0
        else:
            assert not bool(torch.isfinite(lp[0]))            # loud, not wrong
