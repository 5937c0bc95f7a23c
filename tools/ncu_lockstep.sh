M="dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,lts__t_sector_hit_rate.pct"
F="--steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-torch-baseline --no-recompute --no-parity --no-other-configs"
for mode in on off; do
  if [ $mode = off ]; then export B200GRPO_TUNABLES=lockstep_fwd=0,lockstep_dh=0,lockstep_dw=0; else unset B200GRPO_TUNABLES; fi
  ncu --metrics $M --clock-control none -k regex:gemm_sm100 --launch-skip 9 -c 3 --csv --log-file gpurun_out/ncu_lockstep_$mode.csv python bench.py $F > gpurun_out/ncu_lockstep_$mode.log 2>&1
  echo $mode rc=$?
done
