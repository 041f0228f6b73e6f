set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 < /dev/null | tail -2
python bench.py --steps 50 --warmup 10 > gpurun_out/r1i_bench.json 2> gpurun_out/r1i_bench.err < /dev/null; tail -2 gpurun_out/r1i_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1i_bench_ref.json 2> gpurun_out/r1i_bench_ref.err < /dev/null; tail -2 gpurun_out/r1i_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1i_launches.csv python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline > gpurun_out/r1i_ncu_launch.log 2>&1 < /dev/null
bash scripts/ncu_capture.sh r1i < /dev/null
for c in cfg3 cfg5 cfg1; do python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r1i_$c.json 2> gpurun_out/r1i_$c.err < /dev/null; done
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 < /dev/null | tail -2
ls gpurun_out/r1i_* | wc -l
