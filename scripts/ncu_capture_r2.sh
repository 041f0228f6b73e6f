#!/bin/bash
# Round 2: one `ncu --set full` capture per hot kernel of the cfg2 training step (run on the GPU box, AFTER the same bench.py
# command exited 0 without ncu).  Usage: scripts/ncu_capture_r2.sh <tag>   -> gpurun_out/<tag>_<kernel>.ncu-rep
# Template instances are told apart on the DEMANGLED name (--kernel-name-base demangled; the default base is the bare
# function name, on which a regex naming template arguments matches nothing).
tag=${1:-r2}
only=${2:-}          # optional: capture only this name
run() {  # name regex skip count [env]
  if [ -n "$only" ] && [[ " $only " != *" $1 "* ]]; then return; fi
  env $5 ncu --set full --clock-control none --import-source on --kernel-name-base demangled --kernel-name "regex:$2" \
      --launch-skip "$3" --launch-count "$4" -f -o "gpurun_out/${tag}_$1" \
      python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline > "gpurun_out/${tag}_$1.log" 2>&1
}
run attn_bwd 'attn_bwd_stream_kernel' 20 1
run attn_fwd 'attn_fwd_stream_kernel' 20 1
run pair 'gemm_ps2_kernel' 4 2                                     # the CTA-pair (cta_group::2) instances of one step
run ce_stats 'gemm_ps2?_kernel<.*\(int\)1>' 2 1                    # EPI 1: vocabulary projection + online-softmax partials
run ce_grad 'gemm_ps2?_kernel<.*\(int\)2>' 2 1                     # EPI 2: recompute + d logits (bf16 hi/lo copies only)
run loop 'attn_loop_fwd_kernel' 2 1 ALEXCAP_TIME_LOOP=1            # the persistent forward time loop (opt-in)
run gemm_step 'gemm_tc_kernel' 20 3                                # three per-step split-K GEMMs of the time loops (query, gates, ...)
