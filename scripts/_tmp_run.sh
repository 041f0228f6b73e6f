timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_kernels.py tests/test_gpu_gemm_tc.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python scripts/err_report.py 2>&1 | tail -21
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r2q_bench.json 2> gpurun_out/r2q_bench.err; tail -3 gpurun_out/r2q_bench.err; python - <<'PY'
import json
d = json.load(open('gpurun_out/r2q_bench.json'))
print(d['value'], d['ms_per_step'], d['e2e'], d['roofline']['ms_per_launch'], d['roofline']['backward'])
PY
