#!/bin/bash
# One `ncu --set full` capture per hot kernel of the cfg2 training step (run on the GPU box, AFTER bench.py exited 0
# without ncu).  Usage: scripts/ncu_capture.sh <tag>   -> gpurun_out/<tag>_<kernel>.ncu-rep
tag=${1:-r1}
run() {  # name regex skip count
  ncu --set full --clock-control none --import-source on --kernel-name "regex:$2" --launch-skip "$3" --launch-count "$4" -f \
      -o "gpurun_out/${tag}_$1" python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline > "gpurun_out/${tag}_$1.log" 2>&1
}
run attn_fwd 'attn_fwd_stream_kernel' 20 1
run attn_bwd 'attn_bwd_stream_kernel' 20 1
run gemm_step 'gemm_tc_kernel' 20 2      # the two per-step forward GEMMs (query, gates)
run gemm_ps 'gemm_ps_kernel' 12 5        # one step's big GEMMs: Ws projection, embedding->gates, vocabulary projection, dW_o
run attn_dws 'attn_dws_kernel' 1 1
