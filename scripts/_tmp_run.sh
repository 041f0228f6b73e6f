python -m pytest tests/test_gpu_gemm_tc.py -m gpu -x -q -k "bf16" 2>&1 | tail -12
python -m pytest tests/test_gpu_train.py -m gpu -x -q -k "bf16" 2>&1 | tail -12
