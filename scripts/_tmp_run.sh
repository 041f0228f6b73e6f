timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -5
timeout 300 python scripts/err_report.py 2>&1 | tail -50
python bench.py --steps 30 --warmup 5 > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err; tail -3 gpurun_out/r2l_bench.err; python - <<'PY'
import json
d = json.load(open('gpurun_out/r2l_bench.json'))
print(d['value'], d['ms_per_step'], d['e2e'], d['gpu_launches_per_step'], d['roofline_gemm'])
PY
