timeout 600 python -m pytest tests/test_gpu_train.py tests/test_gpu_kernels.py -m gpu -x -q 2>&1 < /dev/null | tail -2
for c in 2 4 6 8; do echo "ctas/sm $c: $(ALEXCAP_DWS_CTAS_PER_SM=$c python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null < /dev/null | cut -c130-175)"; done
