#!/bin/bash
# DRAM traffic of the dW GEMM (ncu, one launch) for a list of tunable sets, e.g.
#   tools/dw_traffic_sweep.sh "dw_group_n=0" "hint_resident=2" "hint_resident=2 dw_group_m=2"
mkdir -p gpurun_out
for cfg in "$@"; do
  args=""; for kv in $cfg; do args="$args --tunable $kv"; done
  tag=$(echo $cfg | tr ' =' '__')
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-torch-baseline $args > /dev/null 2>&1 && \
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:gemm_sm100 --launch-skip 3 -c 3 --csv --log-file gpurun_out/dw_sweep_$tag.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-torch-baseline $args > /dev/null 2>&1
  echo "== $cfg"
  for k in "0, 0, 1, 2, 256, 4" "0, 1, 1, 2, 256, 6" "1, 0, 0, 2, 256, 6"; do
    grep "$k" gpurun_out/dw_sweep_$tag.csv | awk -F'","' -v k="$k" '{printf "%s %s %s | ", k, $(NF-2), $(NF)}' | tr -d '"'; echo
  done
done
