timeout 600 python -m pytest tests/test_gpu_train.py -m gpu -x -q -k "nchw" 2>&1 < /dev/null | grep -E "^E  |passed|failed" | cut -c1-250 | head
