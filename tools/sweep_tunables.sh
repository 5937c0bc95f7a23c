#!/bin/bash
# Same-box sweep of library tunables on the C3 step (or SWEEP_FLAGS="--config c4").  Each argument is one B200GRPO_TUNABLES string
# ("name=value,name=value"); the shipped configuration is re-run every 4 settings so that box drift shows.
#   tools/sweep_tunables.sh fwd_group_m=32 hint_resident=2,hint_stream=1 ...     prints: setting  ms/step  fwd dH dW ms  SM MHz
F="--steps 48 --warmup 3 --no-cpu-baseline --no-torch-baseline --no-e2e --no-recompute --no-parity --no-other-configs ${SWEEP_FLAGS}"
run() {
  python bench.py $F 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels_ms_per_step']; print('$1', round(d['ms_per_step'],3), k['fwd_gemm_stats'], k['bwd_dH_gemm'], k['bwd_dW_gemm'], d['clocks']['sm_mhz'])"
}
i=0
for s in "$@"; do
  if [ $((i % 4)) = 0 ]; then unset B200GRPO_TUNABLES; run shipped; fi
  export B200GRPO_TUNABLES="$s"
  run "$s"
  i=$((i+1))
done
unset B200GRPO_TUNABLES; run shipped
