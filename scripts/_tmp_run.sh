set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python -m pytest tests/test_gpu_decode.py -m gpu -x -q -s -k "config4 and bf16x3" 2>&1 | grep cfg4
python bench.py --steps 30 --warmup 5 > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err; tail -3 gpurun_out/r2i_bench.err; cat gpurun_out/r2i_bench.json
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
