"""Generates the data-side golden fixtures.  Run in the build container only:

    python tests/golden/make_data_golden.py

1. writes a tiny dataset in the reference's on-disk format (`face2text_tiny.h5` + `face2text_tiny.json`, the datasets and
   dictionary keys of AlexCap/my_model_preprocess.py:296-329) with imagecaptioning_b200.h5lite.write;
2. runs the UNMODIFIED reference loader (`/root/reference/AlexCap/MyDataLoader.py`) over it and records what its
   `get_batch` returns for a fixed call sequence -> `data_batches.npz`.

h5py and matplotlib are not installed here: `h5py.File` is bound to `h5lite.File` for this process (the reference only
uses `File(path, 'r')`, `f[name][...]`, `.shape`), and `matplotlib.pyplot` (imported by AlexCap/my_utils.py, unused by
the loader) to an empty module.  The reader itself is pinned on a libhdf5-written file (tests/test_data.py).
"""
import json
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from imagecaptioning_b200 import h5lite  # noqa: E402

CALLS = [({'split': 0, 'iterate': True}, 4, -1)] * 9 + [({'split': 1, 'iterate': True}, 1, -1)] * 6 + \
        [({'split': 2, 'iterate': True}, 2, -1)] * 3 + [({'split': 0, 'iterate': False}, 5, -1)] * 3 + \
        [({'split': 2, 'iterate': False}, 1, 2)]


def make_files():
    rng = np.random.default_rng(123)
    M, T, V = 41, 16, 50
    lengths = rng.integers(3, T + 1, M)
    labels = np.zeros((M, T), np.int32)
    for i, n in enumerate(lengths):
        labels[i, :n] = rng.integers(1, V + 1, n)
    split = np.zeros(M, dtype=int)
    split[30:36] = 1
    split[36:] = 2
    arrays = {
        'split': split, 'attributes': rng.integers(-1, 2, (M, 40)).astype(np.int64), 'lengths': lengths.astype(np.int64),
        'labels': labels, 'img_to_first_phr': np.arange(M, dtype=np.int32), 'img_to_last_phr': np.arange(M, dtype=np.int32),
        'images': rng.integers(0, 256, (M, 12, 10, 3)).astype(np.uint8),
        'features': np.maximum(rng.standard_normal((M, 6, 32)), 0).astype(np.float32),
    }
    info = {
        'token_to_idx': {f'w{i}': i for i in range(1, V + 1)}, 'idx_to_token': {str(i): f'w{i}' for i in range(1, V + 1)},
        'filename_to_idx': {f'{i:05d}.jpg': i for i in range(M)}, 'idx_to_filename': {str(i): f'{i:05d}.jpg' for i in range(M)},
        'attributes_labels': [f'attr{i}' for i in range(40)],
    }
    h5lite.write(os.path.join(HERE, 'face2text_tiny.h5'), arrays)
    with open(os.path.join(HERE, 'face2text_tiny.json'), 'w') as f:
        json.dump(info, f)


def reference_loader():
    h5 = types.ModuleType('h5py')
    h5.File = h5lite.File
    sys.modules['h5py'] = h5
    mpl, plt = types.ModuleType('matplotlib'), types.ModuleType('matplotlib.pyplot')
    mpl.pyplot = plt
    sys.modules.setdefault('matplotlib', mpl)
    sys.modules.setdefault('matplotlib.pyplot', plt)
    sys.path.insert(0, '/root/reference')
    from AlexCap.MyDataLoader import AlexDataLoader
    return AlexDataLoader({'data_h5': os.path.join(HERE, 'face2text_tiny.h5'),
                           'data_json': os.path.join(HERE, 'face2text_tiny.json'), 'device': 'cpu'})


def run_calls(loader):
    torch.manual_seed(123)
    out = {}
    for n, (opt, bs, idx) in enumerate(CALLS):
        img, lab, info, attr = loader.get_batch(opt, bs, idx)
        out[f'img_{n}'] = img.numpy().astype(np.float32).mean(axis=(2, 3)) if img.dim() == 4 else img.numpy()
        out[f'lab_{n}'] = lab.numpy()
        out[f'attr_{n}'] = attr.numpy()
        out[f'files_{n}'] = np.array(info[0]['filename'])
        out[f'bounds_{n}'] = np.array([np.max(info[0]['split_bounds'][0]), info[0]['split_bounds'][1]])
        out[f'shape_{n}'] = np.array(img.shape)
    return out


if __name__ == '__main__':
    make_files()
    ref = reference_loader()
    out = run_calls(ref)
    out['meta'] = np.array([ref.num_images, ref.num_channels, ref.seq_length, ref.vocab_size, len(ref.train_ix),
                            len(ref.val_ix), len(ref.test_ix)])
    np.savez_compressed(os.path.join(HERE, 'data_batches.npz'), **out)
    print('wrote', len(out), 'arrays; first batch files', out['files_0'])
