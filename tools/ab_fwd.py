#!/usr/bin/env python
"""Same-box A/B of two builds of the library on the forward kernel (K1, saved mode) at the C3 shape.

    B200GRPO_LIB_PATH=/path/to/other/libb200grpo.so python tools/ab_fwd.py      (run alternately with the in-tree build)

Prints one line: ms per forward (GEMM + merge, CUDA events over 40 launches after 5 warm-ups) and the SM clock seen."""
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pipelinerl_swe_b200 import _abi, ops  # noqa: E402

d = torch.device("cuda:0")
M, H, V = 8191, 3584, 152064
g = torch.Generator(device=d).manual_seed(1)
h = torch.randn(M, H, device=d, generator=g).bfloat16()
w = (torch.randn(V, H, device=d, generator=g) * 0.02).bfloat16()
tgt = torch.randint(0, V, (M,), device=d, generator=g)
saved = ops.lmhead_saved_buffer(M, H, V, d)
for _ in range(5):
    ops.lmhead_fwd_stats(h, w, tgt, 1.0, saved=saved)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(40):
    ops.lmhead_fwd_stats(h, w, tgt, 1.0, saved=saved)
e1.record()
torch.cuda.synchronize()
clk = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
print(f"{os.path.basename(os.path.dirname(_abi.LIB_PATH))}/{os.path.basename(_abi.LIB_PATH)} abi={_abi.lib().b200grpo_abi_version()} fwd {e0.elapsed_time(e1) / 40:.3f} ms  [{clk}]")
