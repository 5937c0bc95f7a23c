"""Campaign runner of the kernel-level lm_head fuzz (tests/gpu_probe.py probe_lmhead_fuzz: K1 / finalize / K2 through the C ABI against
torch autograd in fp32 over random shapes, tile variants, tunables, shards, hot and confident targets).

    python tests/fuzz_lmhead.py <first_seed> <n_seeds>      (GPU box)
"""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_probe import probe_lmhead_fuzz
nbad = 0
for seed in range(int(sys.argv[1]), int(sys.argv[1]) + int(sys.argv[2])):
    ok, msg = probe_lmhead_fuzz(seed, 16)
    nbad += not ok
    print("LMHEAD FUZZ", "OK" if ok else "FAIL", seed, msg[:700], flush=True)
print("SUMMARY failing seeds:", nbad)
