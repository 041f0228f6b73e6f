timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_decode.py -m gpu -x -q 2>&1 < /dev/null | tail -3
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r3k_bench.json 2> gpurun_out/r3k_bench.err < /dev/null; tail -3 gpurun_out/r3k_bench.err; cut -c50-200 gpurun_out/r3k_bench.json
