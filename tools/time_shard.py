#!/usr/bin/env python
"""K1 + K2 of one vocab shard (config 4: H=5120, Vloc=19008 of V=152064, T=8192) on one GPU under different soft-lockstep settings.

    python tools/time_shard.py [Vloc] [H]     prints ms per forward+backward for each setting (same process, interleaved twice)
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pipelinerl_swe_b200 import _abi, ops  # noqa: E402

Vloc = int(sys.argv[1]) if len(sys.argv) > 1 else 19008
H = int(sys.argv[2]) if len(sys.argv) > 2 else 5120
M = 8191
d = torch.device("cuda:0")
g = torch.Generator(device=d).manual_seed(1)
h = torch.randn(M, H, device=d, generator=g).bfloat16()
w = (torch.randn(Vloc, H, device=d, generator=g) * 0.02).bfloat16()
tgt = torch.randint(0, 152064, (M,), device=d, generator=g)
c_lp = torch.randn(M, device=d, generator=g) * 0.1
gw = torch.zeros(Vloc, H, device=d)
names = ("lockstep_fwd", "lockstep_dh", "lockstep_dw", "lockstep_slack")
shipped = tuple(_abi.set_tunable(n, 0) for n in names)


def step():
    saved = ops.lmhead_saved_buffer(M, H, Vloc, d)
    stats = ops.lmhead_fwd_stats(h, w, tgt, 1.0, v_offset=0, saved=saved)
    lp, ent, lse = ops.lmhead_finalize(stats)
    ops.lmhead_bwd(h, w, tgt, lse, ent, c_lp, None, 1.0, v_offset=0, grad_weight_out=gw, accumulate=True, saved=saved)


def timed(setting, n=40):
    for nm, v in zip(names, setting):
        _abi.set_tunable(nm, v)
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


settings = [shipped, (0, 0, 0, 2), (shipped[0], shipped[1], 0, 2), (0, shipped[1], 0, 2), (8, 8, 0, 2), (16, 16, 64, 2)]
for rnd in range(2):
    for s in settings:
        print(f"Vloc={Vloc} H={H} lockstep fwd,dh,dw,slack={s}: {timed(s):.3f} ms", flush=True)
for nm, v in zip(names, shipped):
    _abi.set_tunable(nm, v)
