set -x
python bench.py --steps 50 --warmup 10 > gpurun_out/r1g_bench.json 2> gpurun_out/r1g_bench.err; tail -2 gpurun_out/r1g_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1g_bench_ref.json 2> gpurun_out/r1g_bench_ref.err; tail -2 gpurun_out/r1g_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1g_launches.csv python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline > gpurun_out/r1g_ncu_launch.log 2>&1
bash scripts/ncu_capture.sh r1g
ls -la gpurun_out/r1g_*
