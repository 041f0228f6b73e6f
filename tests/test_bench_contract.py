"""CPU: the parts of bench.py's driver contract that need no GPU -- the reference arm's JSON line (the reference's CPU
path timed through oracle/ref_port.py), its behaviour under a torchrun-style environment (rank 0 alone works), and that
the GPU arm refuses to run without CUDA instead of falling back."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), *args], capture_output=True, text=True, env=e,
                          timeout=600)


def test_reference_arm_json_line():
    r = _run(['--impl', 'reference', '--config', 'cfg1', '--steps', '2', '--warmup', '1'])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1                                  # ONE JSON line on stdout
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and d['unit'] == 'captions/s' and d['higher_is_better'] is True
    assert d['metric'].startswith('captions/sec') and d['n_gpus'] == 1 and d['steps'] == 2 and d['warmup'] == 1
    assert d['value'] > 0 and d['ms_per_step'] > 0 and d['vs_baseline'] is None and d['scaling'] == 'weak'
    assert d['value'] == pytest.approx(16 / (d['ms_per_step'] * 1e-3), rel=1e-3)          # captions of one batch / step time
    assert d['config']['workload'].startswith('LSTMModel') and 'model' not in d['config']
    cb = d['cpu_baseline']
    assert cb['kind'] == 'port' and cb['value'] == d['value'] and cb['cores'] >= 1 and cb['sample']
    assert d['e2e'] == {'value': d['value'], 'unit': 'captions/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}


def test_reference_arm_other_ranks_exit_without_work():
    r = _run(['--impl', 'reference', '--config', 'cfg1', '--gpus', '2', '--steps', '1', '--warmup', '0'],
             env={'RANK': '1', 'LOCAL_RANK': '1', 'WORLD_SIZE': '2'})
    assert r.returncode == 0 and r.stdout.strip() == ''


@pytest.mark.skipif(torch.cuda.is_available(), reason='checks the no-GPU behaviour')
def test_gpu_arm_has_no_cpu_fallback():
    r = _run(['--config', 'cfg1', '--steps', '1', '--warmup', '3'])
    assert r.returncode != 0 and r.stdout.strip() == ''
    assert 'no CPU fallback' in r.stderr
