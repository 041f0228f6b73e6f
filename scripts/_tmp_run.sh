timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_train.py -m gpu -x -q 2>&1 < /dev/null | tail -3
timeout 120 python scripts/bench_attn.py 2>&1 < /dev/null | tail -1 | sed 's/.*} //'
timeout 300 python scripts/err_report.py 2>&1 < /dev/null | tail -21 | sed -n 9,16p
python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null < /dev/null | cut -c50-200
