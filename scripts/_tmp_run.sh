python -m pytest tests/test_gpu_train.py -m gpu -x -q -k "two_phase" 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r1v_n1.json 2> gpurun_out/r1v_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r1v_n2.json 2> gpurun_out/r1v_n2.err
tail -3 gpurun_out/r1v_n2.err
