"""ms/step of the resident cfg2 step under the current environment (timing experiments): python scripts/quick_ms.py [cfg]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
c = bench.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else 'cfg2']
run = bench.TrainRun(c, torch.device('cuda', 0), 0, 1, None, graph=True)
ms, _ = run.time_resident(20, 5)
print('%s ms_per_step %.4f loss %.5f launches %d' % (os.environ.get('TAG', ''), ms, run.loss, run.launches_per_step()))
