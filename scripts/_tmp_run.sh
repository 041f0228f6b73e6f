N=$1
s=$(date +%s)
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 50 --warmup 10 > gpurun_out/r2y_n$N.json 2> gpurun_out/r2y_n$N.err < /dev/null
echo "exit code $? after $(( $(date +%s) - s )) s"
tail -3 gpurun_out/r2y_n$N.err; cut -c50-260 gpurun_out/r2y_n$N.json; grep -o '"e2e": {[^}]*}' gpurun_out/r2y_n$N.json
