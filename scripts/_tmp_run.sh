N=$1
s=$(date +%s)
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 50 --warmup 10 > gpurun_out/r1j_n$N.json 2> gpurun_out/r1j_n$N.err < /dev/null
echo "exit code $? after $(( $(date +%s) - s )) s"
cut -c50-260 gpurun_out/r1j_n$N.json
