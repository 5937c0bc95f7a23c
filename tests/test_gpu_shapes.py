"""K1 / K2 parity AT THE BENCHMARKED SHAPES, default tunables (the production tile schedule): BASELINE config 3
(Qwen2.5-7B: M=8191, H=3584, V=152064) and config 4 (Qwen2.5-32B: H=5120; full vocab and the V/8 shard), in `saved` and
`recompute` backward modes, against the reference's fp32 math (rl/__init__.py:197-210 + autograd) on the same bf16 values:
every row's logprob / entropy (1e-3 relative), the FULL dH and the FULL dW (2e-2 of max) -- so every tile of the
schedule is covered, the tail-split tiles and the ragged 152064 % 256 vocab tail included -- with the gradient
accumulated over two steps into a non-zero f32 buffer (the `red.global.add.v4.f32` epilogue)."""
import pytest
import torch

import shape_parity as sp

pytestmark = pytest.mark.gpu

SHAPES = {
    # name: (M, H, V, temperature)
    "c3": (8191, 3584, 152064, 1.0),
    "c4": (8191, 5120, 152064, 0.7),
}
_cache: dict = {}


def _problem(name):
    """Seeded inputs + the fp32 reference of one shape (computed once per session; ~5 GB on the device)."""
    if name in _cache:
        return _cache[name]
    _cache.clear()   # one shape resident at a time
    torch.cuda.empty_cache()
    M, H, V, temp = SHAPES[name]
    d = torch.device("cuda:0")
    g = torch.Generator(device=d).manual_seed(len(name) + H)
    hidden = torch.randn(M, H, device=d, generator=g).bfloat16()
    weight = (torch.randn(V, H, device=d, generator=g) * 0.02).bfloat16()
    targets = torch.randint(0, V, (M,), device=d, generator=g)
    targets[:64] = torch.arange(V - 64, V, device=d)            # targets inside the ragged last tile
    labelled = (torch.rand(M, device=d, generator=g) < 0.7).float()
    c_lp = torch.randn(M, device=d, generator=g) * 0.1 * labelled   # masked positions carry a zero coefficient, as from K3
    c_ent = torch.randn(M, device=d, generator=g) * 0.01 * labelled
    p = dict(hidden=hidden, weight=weight, targets=targets, inv_temp=1.0 / temp, c_lp=c_lp, c_ent=c_ent)
    p["ref"] = sp.reference_lmhead(hidden, weight, targets, 1.0 / temp, c_lp, c_ent)
    _cache[name] = p
    return p


@pytest.mark.parametrize("name", ["c3", "c4"])
def test_fast_reference_is_the_fp32_reference(name):
    """The chunked comparator (bf16 x bf16 products accumulated in fp32) equals plain fp32 matmul + log_softmax."""
    p = _problem(name)
    assert sp.fast_reference_is_fp32(p["hidden"], p["weight"], p["targets"], p["inv_temp"]) < 2e-4   # |logprob| ~ 12: 2e-5 relative, the summation-order noise of the plain fp32 matmul itself


@pytest.mark.parametrize("mode", ["saved", "saved_no_entropy", "recompute"])
@pytest.mark.parametrize("name", ["c3", "c4"])
def test_lmhead_fwd_bwd_at_benchmarked_shape(name, mode):
    """saved: entropy coefficients present -> dlogits pass on the saved exponentials + two GEMMs; saved_no_entropy: what the
    benchmarked config runs (entropy_bonus == 0) -> the GEMMs read the saved exponentials directly (row-scale epilogue in dH,
    pre-scaled hidden + sparse target rows in dW); recompute: third GEMM."""
    p = _problem(name)
    if mode == "saved_no_entropy":
        if "ref0" not in p:
            p["ref0"] = sp.reference_lmhead(p["hidden"], p["weight"], p["targets"], p["inv_temp"], p["c_lp"], torch.zeros_like(p["c_lp"]))
        rec = sp.check_lmhead_at_shape(p["hidden"], p["weight"], p["targets"], p["inv_temp"], p["c_lp"], None, save=True, ref=p["ref0"])
    else:
        rec = sp.check_lmhead_at_shape(p["hidden"], p["weight"], p["targets"], p["inv_temp"], p["c_lp"], p["c_ent"],
                                       save=(mode == "saved"), ref=p["ref"])
    print(name, mode, rec)
    assert rec["lp_rel"] < sp.LP_RTOL and rec["ent_rel"] < sp.LP_RTOL, rec
    assert rec["dH_rel"] < sp.GRAD_RTOL and rec["dW_rel"] < sp.GRAD_RTOL, rec
    assert rec["ok"], rec


@pytest.mark.parametrize("mode", ["saved", "saved_no_entropy", "recompute"])
@pytest.mark.parametrize("rank", [0, 7])
def test_lmhead_c4_vocab_shard_of_eight(rank, mode):
    """Config 4's per-rank problem (Vloc = 19008 rows at v_offset = rank * 19008): shard dW and the dH partial against the
    full-vocab reference restricted to those rows."""
    from pipelinerl_swe_b200.vocab_parallel import shard_range
    p = _problem("c4")
    V = p["weight"].shape[0]
    lo, hi = shard_range(V, 8, rank)
    c_ent = None if mode == "saved_no_entropy" else p["c_ent"]
    ref = sp.reference_lmhead(p["hidden"], p["weight"], p["targets"], p["inv_temp"], p["c_lp"],
                              c_ent if c_ent is not None else torch.zeros_like(p["c_lp"]), lo, hi)
    rec = sp.check_lmhead_at_shape(p["hidden"], p["weight"], p["targets"], p["inv_temp"], p["c_lp"], c_ent,
                                   save=(mode != "recompute"), v_lo=lo, v_hi=hi, ref=ref)
    print("c4 shard", rank, mode, rec)
    assert rec["ok"], rec


def test_lmhead_c4_all_eight_shards_merge_to_the_unsharded_logprobs():
    """finalize over the 8 shard partials (rank order) at the C4 shape == the fp32 reference over the whole vocab."""
    from pipelinerl_swe_b200 import ops
    from pipelinerl_swe_b200.vocab_parallel import shard_range
    p = _problem("c4")
    V = p["weight"].shape[0]
    parts = []
    for r in range(8):
        lo, hi = shard_range(V, 8, r)
        parts.append(ops.lmhead_fwd_stats(p["hidden"], p["weight"][lo:hi].contiguous(), p["targets"], p["inv_temp"], v_offset=lo))
    lp, ent, lse = ops.lmhead_finalize(torch.stack(parts))
    assert sp.rel_max(lp, p["ref"]["lp"]) < sp.LP_RTOL
    assert sp.rel_max(ent, p["ref"]["ent"]) < sp.LP_RTOL
    assert sp.rel_max(lse, p["ref"]["lse"]) < sp.LP_RTOL
