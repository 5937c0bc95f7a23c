"""Parity of K1 / K2 at the shapes that are benchmarked (C3: H=3584, C4: H=5120, V=152064, T=8192).

Test infrastructure (used by tests/test_gpu_shapes.py and by bench.py's pre-run parity gate), never by the product.

The comparator is the reference's own math -- ``logits / T -> log_softmax -> softmax -> entropy -> gather`` and torch
autograd through it (/root/reference/pipelinerl/finetune/rl/__init__.py:197-210) -- in fp32 on the same bf16 values,
row-chunked on the GPU so that the [rows x vocab] fp32 logits fit.  Two things make the full shape affordable:

* the product of two bf16 numbers is exact in fp32, so ``torch.mm(bf16, bf16, out_dtype=float32)`` (fp32 accumulate on
  the tensor cores) IS the fp32 matmul of those values up to summation order; ``fast_reference_is_fp32`` proves that
  against a plain fp32 ``@`` (TF32 off) on a row sample;
* the fp32 dlogits of the reference are split into hi + lo bf16 terms (relative error 2^-17) for the two gradient
  GEMMs, so the FULL dH [M,H] and the FULL dW [V,H] are compared, every tile of the production schedule included
  (tail-split tiles, the ragged last vocab tile, zig-zag K order).
"""
from __future__ import annotations

import torch

LP_RTOL = 1e-3     # north star: logprobs / loss within 1e-3 relative
GRAD_RTOL = 2e-2   # gradients within 2e-2 relative (of the largest reference entry)


def _mm32(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """fp32-accumulated product of two bf16 matrices (exact products)."""
    return torch.mm(a, b, out_dtype=torch.float32)


def _split_bf16(x: torch.Tensor):
    hi = x.to(torch.bfloat16)
    lo = (x - hi.float()).to(torch.bfloat16)
    return hi, lo


def reference_lmhead(hidden, weight, targets, inv_temp, c_lp, c_ent, v_lo=0, v_hi=None, rows=1024, want_grads=True):
    """fp32 reference over the FULL vocab ``weight`` [V,H]; gradients restricted to the vocab rows [v_lo, v_hi) (a shard).

    Returns dict(lp, ent, lse [M] f32; dH [M,H] f32 = contribution of the rows [v_lo, v_hi); dW [v_hi-v_lo, H] f32)."""
    M, H = hidden.shape
    V = weight.shape[0]
    v_hi = V if v_hi is None else v_hi
    dev = hidden.device
    lp = torch.empty(M, dtype=torch.float32, device=dev)
    ent = torch.empty_like(lp)
    lse = torch.empty_like(lp)
    dH = torch.zeros(M, H, dtype=torch.float32, device=dev) if want_grads else None
    dW = torch.zeros(v_hi - v_lo, H, dtype=torch.float32, device=dev) if want_grads else None
    wt = weight.t()
    w_sh = weight[v_lo:v_hi]
    for r0 in range(0, M, rows):
        r1 = min(M, r0 + rows)
        z = _mm32(hidden[r0:r1], wt).requires_grad_(want_grads)             # the model's logits (:197)
        logits = z * inv_temp                                               # / temperature (:198-199)
        logprobs = torch.log_softmax(logits, dim=-1)                        # :202
        probs = torch.softmax(logits, dim=-1)                               # :203
        entropy = -(probs * logprobs).sum(dim=-1)                           # :204
        new_lp = torch.gather(logprobs, 1, targets[r0:r1, None]).squeeze(1)  # :206-208
        lp[r0:r1], ent[r0:r1] = new_lp.detach(), entropy.detach()
        lse[r0:r1] = torch.logsumexp(logits.detach(), dim=-1)
        if want_grads:
            (dz,) = torch.autograd.grad((new_lp * c_lp[r0:r1]).sum() + (entropy * c_ent[r0:r1]).sum(), z)
            dz = dz[:, v_lo:v_hi]
            hi, lo = _split_bf16(dz)
            dH[r0:r1] = _mm32(hi, w_sh) + _mm32(lo, w_sh)
            hc = hidden[r0:r1]
            dW += _mm32(hi.t(), hc) + _mm32(lo.t(), hc)
        del z, logits, logprobs, probs
    return dict(lp=lp, ent=ent, lse=lse, dH=dH, dW=dW)


def fast_reference_is_fp32(hidden, weight, targets, inv_temp, n_rows=192) -> float:
    """max |logprob| difference between the chunked fast reference and a plain fp32 ``@`` (TF32 off) on a row sample."""
    M = hidden.shape[0]
    idx = torch.linspace(0, M - 1, n_rows, device=hidden.device).long()
    zero = torch.zeros(n_rows, device=hidden.device)
    fast = reference_lmhead(hidden[idx].contiguous(), weight, targets[idx], inv_temp, zero, zero, want_grads=False)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        z = (hidden[idx].float() @ weight.float().t()) * inv_temp
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    logp = torch.log_softmax(z, -1)
    lp = logp.gather(1, targets[idx, None]).squeeze(1)
    ent = -(logp.exp() * logp).sum(-1)
    return max(float((fast["lp"] - lp).abs().max()), float((fast["ent"] - ent).abs().max()))


def rel_max(got: torch.Tensor, want: torch.Tensor) -> float:
    return float((got.float() - want).abs().max() / (want.abs().max() + 1e-30))


def tail_tile_rows(n_rows: int, n_cols: int, tile_rows: int, clusters: int = 74):
    """Row ranges of the output covered by the tiles of the last partial wave (the ones the tail split cuts along K).
    Tiles are 256 columns wide; with group_m = 1 rasterisation the last tiles % clusters tiles are the last ones in
    row-block-major order, i.e. the last row blocks of the output."""
    m_blocks, n_blocks = -(-n_rows // tile_rows), -(-n_cols // 256)
    r = (m_blocks * n_blocks) % clusters
    if r == 0:
        return None
    first_blk = (m_blocks * n_blocks - r) // n_blocks
    return first_blk * tile_rows, n_rows


def check_lmhead_at_shape(hidden, weight_full, targets, inv_temp, c_lp, c_ent, *, save: bool, v_lo=0, v_hi=None,
                          accumulate_steps=2, ref=None) -> dict:
    """Run K1 -> finalize -> K2 of this library with the DEFAULT tunables on (a vocab shard of) the given problem and
    compare with ``reference_lmhead``.  ``accumulate_steps`` backward passes add into a non-zero f32 gradient
    (``red.global.add`` path).  ``c_ent=None``: the entropy-free backward (GEMMs straight on the saved exponentials); ``ref``
    must then come from zero entropy coefficients.  Returns the error record; ``ok`` applies the north-star tolerances."""
    from pipelinerl_swe_b200 import ops
    M, H = hidden.shape
    V = weight_full.shape[0]
    v_hi = V if v_hi is None else v_hi
    sharded = (v_lo, v_hi) != (0, V)
    dev = hidden.device
    w = weight_full[v_lo:v_hi] if not sharded else weight_full[v_lo:v_hi].contiguous()
    if ref is None:
        ref = reference_lmhead(hidden, weight_full, targets, inv_temp, c_lp, c_ent if c_ent is not None else torch.zeros_like(c_lp),
                               v_lo, v_hi)
    g = torch.Generator(device=dev).manual_seed(17)
    scale = float(ref["dW"].abs().max())
    g0 = torch.randn(v_hi - v_lo, H, device=dev, generator=g) * scale      # the gradient already in .grad
    gw = g0.clone()
    out = {}
    for step in range(accumulate_steps):
        saved = ops.lmhead_saved_buffer(M, H, v_hi - v_lo, dev) if save else None
        stats = ops.lmhead_fwd_stats(hidden, w, targets, inv_temp, v_offset=v_lo, saved=saved)
        if sharded:   # the other shards' partials come from the reference side: only (lse, entropy) feed the backward
            lse, ent = ref["lse"], ref["ent"]
        else:
            lp, ent, lse = ops.lmhead_finalize(stats)
            out["lp_rel"] = rel_max(lp, ref["lp"])
            out["ent_rel"] = rel_max(ent, ref["ent"])
            out["lp_abs"] = float((lp - ref["lp"]).abs().max())
        gh, _ = ops.lmhead_bwd(hidden, w, targets, lse, ent, c_lp, c_ent, inv_temp, v_offset=v_lo, grad_weight_out=gw,
                               accumulate=True, saved=saved)
        del saved
    out["dH_rel"] = rel_max(gh, ref["dH"])
    got_dw = (gw - g0) / accumulate_steps
    out["dW_rel"] = rel_max(got_dw, ref["dW"])
    # the regions the production schedule treats specially: tail-split tiles of dW (last row blocks) and the ragged vocab tail
    for tile_rows in (256, 512):
        tr = tail_tile_rows(v_hi - v_lo, H, tile_rows)
        if tr is not None:
            out[f"dW_tail_tiles_rel_{tile_rows}"] = float((got_dw[tr[0]:tr[1]] - ref["dW"][tr[0]:tr[1]]).abs().max() / (ref["dW"].abs().max() + 1e-30))
    out["dW_first_tile_rel"] = float((got_dw[:256] - ref["dW"][:256]).abs().max() / (ref["dW"].abs().max() + 1e-30))
    out["dH_tail_rows_rel"] = float((gh[-512:].float() - ref["dH"][-512:]).abs().max() / (ref["dH"].abs().max() + 1e-30))
    out["accumulate_steps"] = accumulate_steps
    out["mode"] = ("saved" if save else "recompute") + ("" if c_ent is not None else ", entropy-free backward")
    ok = out["dH_rel"] < GRAD_RTOL and out["dW_rel"] < GRAD_RTOL
    if not sharded:
        ok = ok and out["lp_rel"] < LP_RTOL and out["ent_rel"] < LP_RTOL
    out["ok"] = bool(ok)
    return out
