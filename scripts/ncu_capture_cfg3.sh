#!/bin/bash
# cfg3 (bf16 mode, B = 512): launch list of two steps + `ncu --set full` of the bf16-row attention kernels (run on the GPU box,
# AFTER the same bench.py command exited 0 without ncu).  Usage: scripts/ncu_capture_cfg3.sh <tag>
tag=${1:-r2}
cmd="python bench.py --config cfg3 --steps 2 --warmup 3 --no-graph --no-cpu-baseline"
$cmd > gpurun_out/${tag}_cfg3_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1700 --csv --log-file gpurun_out/${tag}_launches_cfg3.csv $cmd > gpurun_out/${tag}_cfg3_list.log 2>&1
for k in attn_fwd attn_bwd; do
  ncu --set full --clock-control none --import-source on --kernel-name "regex:${k}_stream_kernel" --launch-skip 20 --launch-count 1 -f \
      -o gpurun_out/${tag}_cfg3_$k $cmd > gpurun_out/${tag}_cfg3_$k.log 2>&1
done
