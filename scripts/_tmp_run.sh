timeout 600 python -m pytest tests/test_gpu_gemm_tc.py -m gpu -x -q -k "presplit_epilogues" 2>&1 | grep -E "^E  |passed|failed" | cut -c1-250 | head
