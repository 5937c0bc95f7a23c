#!/bin/bash
# Run every bring-up probe in its own process (bounded), collect one line each.
mkdir -p gpurun_out
out=gpurun_out/probes.log
: > $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv >> $out 2>&1
for p in $(python tests/gpu_probe.py list | grep -E "${1:-.}"); do
  timeout 240 python tests/gpu_probe.py $p > gpurun_out/probe_$p.txt 2>&1
  rc=$?
  grep -E "^PROBE|timed out|error|Error" gpurun_out/probe_$p.txt | head -5 >> $out
  [ $rc -ne 0 ] && echo "  (exit $rc) $p" >> $out
done
cat $out
