python bench.py --steps 30 --warmup 5 > gpurun_out/r1f_bench.json 2> gpurun_out/r1f_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1f_bench_ref.json 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r1f_launches.csv python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline > gpurun_out/r1f_ncu_launch.log 2>&1
bash scripts/ncu_capture.sh r1f
