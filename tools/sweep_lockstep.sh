#!/bin/bash
# Same-box sweep of the soft K-lockstep tunables on the C3 step.  Each argument is one setting "fwd,dh,dw[,slack]" (k-blocks per
# epoch, 0 = off); the base configuration (all off) is re-run every 4 settings so that box drift shows.
#   tools/sweep_lockstep.sh 16,16,0 16,8,0 ...        prints: setting  ms/step  fwd dH dW ms  SM MHz
F="--steps 48 --warmup 3 --no-cpu-baseline --no-torch-baseline --no-e2e --no-recompute --no-parity --no-other-configs ${SWEEP_FLAGS}"
run() {
  python bench.py $F 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels_ms_per_step']; print('$1', round(d['ms_per_step'],3), k['fwd_gemm_stats'], k['bwd_dH_gemm'], k['bwd_dW_gemm'], d['clocks']['sm_mhz'])"
}
i=0
for s in "$@"; do
  if [ $((i % 4)) = 0 ]; then unset B200GRPO_TUNABLES; run base; fi
  IFS=, read f h w sl <<< "$s"
  export B200GRPO_TUNABLES="lockstep_fwd=$f,lockstep_dh=$h,lockstep_dw=$w,lockstep_slack=${sl:-2}"
  run "$s"
  i=$((i+1))
done
unset B200GRPO_TUNABLES; run base
