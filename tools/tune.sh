#!/bin/bash
# Run bench.py (C3, device-resident only) once per tunable set; one summary line each.
# usage: tools/tune.sh "hint_resident=2" "hint_resident=2 hint_stream=1" ...
mkdir -p gpurun_out
for cfg in "$@"; do
  args=""; for kv in $cfg; do args="$args --tunable $kv"; done
  python bench.py --steps 24 --warmup 3 --no-e2e --no-cpu-baseline --no-torch-baseline $args 2>/dev/null | python -c "
import sys, json
d = json.loads([l for l in sys.stdin if l.startswith('{')][-1])
k = d['roofline']['kernels_ms_per_step']
print('%-44s %7.3f ms/step %8.0f tok/s  fwd %.2f pass %.2f dH %.2f dW %.2f  clk %s' % ('$cfg', d['ms_per_step'], d['value'], k.get('fwd_gemm_stats', 0), k.get('dlogits_pass', 0), k.get('bwd_dH_gemm', 0), k.get('bwd_dW_gemm', 0), d['clocks']['sm_mhz']))
"
done
