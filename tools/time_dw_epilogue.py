#!/usr/bin/env python
"""dW GEMM at the C3 shape back to back: f32 red.global.add epilogue (accumulate) vs f32 plain store vs bf16 store.
How much of dW is its epilogue traffic?  (saved mode, entropy-free path: the saved exponentials are read-only and reused.)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pipelinerl_swe_b200 import _abi, ops  # noqa: E402

d = torch.device("cuda:0")
M, H, V = 8191, 3584, 152064
g = torch.Generator(device=d).manual_seed(1)
h = torch.randn(M, H, device=d, generator=g).bfloat16()
w = (torch.randn(V, H, device=d, generator=g) * 0.02).bfloat16()
tgt = torch.randint(0, V, (M,), device=d, generator=g)
c_lp = torch.randn(M, device=d, generator=g) * 0.1
saved = ops.lmhead_saved_buffer(M, H, V, d)
stats = ops.lmhead_fwd_stats(h, w, tgt, 1.0, saved=saved)
lp, ent, lse = ops.lmhead_finalize(stats)
gw32 = torch.zeros(V, H, device=d)
gw16 = torch.empty(V, H, device=d, dtype=torch.bfloat16)


def run(kind):
    kw = dict(saved=saved, need_grad_hidden=False, stages=_abi.BWD_DW)
    if kind == "f32_red":
        ops.lmhead_bwd(h, w, tgt, lse, ent, c_lp, None, 1.0, grad_weight_out=gw32, accumulate=True, **kw)
    elif kind == "f32_store":
        ops.lmhead_bwd(h, w, tgt, lse, ent, c_lp, None, 1.0, grad_weight_out=gw32, accumulate=False, **kw)
    else:
        ops.lmhead_bwd(h, w, tgt, lse, ent, c_lp, None, 1.0, grad_weight_out=gw16, accumulate=False, **kw)


# the first call prepares the row scales etc. (stage DLOGITS) once
ops.lmhead_bwd(h, w, tgt, lse, ent, c_lp, None, 1.0, grad_weight_out=gw32, accumulate=True, saved=saved)
for rnd in range(2):
    for kind in ("f32_red", "f32_store", "bf16_store"):
        for _ in range(5):
            run(kind)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(60):
            run(kind)
        e1.record()
        torch.cuda.synchronize()
        print(f"dW {kind}: {e0.elapsed_time(e1) / 60:.3f} ms", flush=True)
