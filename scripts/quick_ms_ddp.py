"""ms/step of the overlapped data-parallel step under the current environment (timing experiments), one rank per GPU:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 scripts/quick_ms_ddp.py [cfg]"""
import os, sys
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
c = bench.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else 'cfg2']
run = bench.TrainRun(c, torch.device('cuda', local), rank, world, dist, graph=True)
ms, _ = run.time_resident(20, 5)
t = torch.tensor([ms], device='cuda')
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print('%s N=%d ms_per_step %.4f' % (os.environ.get('TAG', ''), world, float(t)), flush=True)
dist.barrier()
sys.stdout.flush(); sys.stderr.flush()
os._exit(0)        # (as bench.py: tearing NCCL down after captured collectives can hang)
