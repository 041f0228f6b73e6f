#!/bin/bash
# One `ncu --set full` capture per hot kernel of the cfg2 training step (run on the GPU box, AFTER bench.py exited 0
# without ncu).  Usage: scripts/ncu_capture.sh <tag>   -> gpurun_out/<tag>_<kernel>.ncu-rep
tag=${1:-r1}
run() {  # name regex skip
  ncu --set full --clock-control none --import-source on -k "regex:$2" -s "$3" -c 1 -f -o "gpurun_out/${tag}_$1" \
      python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline > "gpurun_out/${tag}_$1.log" 2>&1
}
run attn_fwd 'attn_fwd_kernel' 20
run attn_bwd 'attn_bwd_kernel' 20
run gemm_step 'gemm_tc_kernel<64, false, false' 40
run gemm_big 'gemm_tc_kernel<128, false, false' 1
run attn_dws 'attn_dws_kernel' 1
