"""CPU: the reference's on-disk formats -- h5lite against a libhdf5-written file, the
writer against the reader, AlexDataLoader against batches recorded from the UNMODIFIED reference loader
(tests/golden/make_data_golden.py) and, in the build container, against that loader run live."""
import os
import sys

import numpy as np
import pytest
import torch

from imagecaptioning_b200 import h5lite
from imagecaptioning_b200.data import AlexDataLoader, load_reference_checkpoint, write_dataset
from tests.conftest import have_reference

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
OPT = {'data_h5': os.path.join(GOLD, 'face2text_tiny.h5'), 'data_json': os.path.join(GOLD, 'face2text_tiny.json'),
       'device': 'cpu'}


def test_reader_on_a_file_written_by_libhdf5():
    """MATLAB v7.3 = HDF5 behind a 512-byte user block, written by the HDF5 library itself (scipy test data:
    `testdouble` = pi/4 * arange(9), scipy/io/matlab/tests/test_mio.py `theta`)."""
    with h5lite.File(os.path.join(GOLD, 'libhdf5_written.mat')) as f:
        assert f.keys() == ['testdouble'] and 'testdouble' in f and '/nope' not in f
        d = f['/testdouble']
        assert d.shape == (9, 1) and d.dtype == np.dtype('<f8')
        assert np.array_equal(d[:].ravel(), np.pi / 4 * np.arange(9))
        with pytest.raises(KeyError):
            f['labels']


@pytest.mark.parametrize('userblock', [0, 512, 2048])
@pytest.mark.parametrize('storage', ['contiguous', 'chunked', 'chunked+deflate', 'chunked+shuffle+deflate'])
def test_writer_reader_round_trip(tmp_path, userblock, storage):
    rng = np.random.default_rng(1)
    ds = {'labels': rng.integers(0, 9000, (37, 16)).astype(np.int32), 'split': np.array([0] * 30 + [1] * 4 + [2] * 3),
          'images': rng.integers(0, 256, (37, 6, 5, 3)).astype(np.uint8), 'attributes': rng.integers(-1, 2, (37, 40)),
          'f32': rng.standard_normal((5, 3)).astype(np.float32), 'f64': rng.standard_normal(11),
          'empty': np.zeros((0, 4), np.int32), 'big_endian': np.arange(5, dtype='>i2'), 'u16': np.arange(9, dtype=np.uint16),
          'scalar_like': np.array([7], np.int8)}
    kw = {}
    if storage != 'contiguous':
        kw = dict(chunks={'labels': (8, 16), 'images': (5, 6, 5, 3), 'f64': (4,)}, deflate='deflate' in storage,
                  shuffle='shuffle' in storage)
    p = str(tmp_path / 'x.h5')
    h5lite.write(p, ds, userblock=userblock, **kw)
    with h5lite.File(p) as f:
        assert f.keys() == sorted(ds)
        for k, v in ds.items():
            a = f[k][:]
            assert a.shape == v.shape and a.dtype == v.dtype and np.array_equal(a, v), k
        ix = [3, 5, 21]
        assert np.array_equal(f['/images'][ix, :, :, :3], ds['images'][ix])          # MyDataLoader.py:82
        assert f['split'][31] == 1 and len(f['labels']) == 37                        # MyDataLoader.py:41
        assert np.array_equal(np.asarray(f['labels']), ds['labels'])


def test_reader_rejects_what_it_does_not_understand(tmp_path):
    p = tmp_path / 'bad.h5'
    p.write_bytes(b'not an hdf5 file' * 64)
    with pytest.raises(h5lite.H5Error):
        h5lite.File(str(p))
    good = str(tmp_path / 'good.h5')
    h5lite.write(good, {'a': np.arange(4)})
    raw = bytearray(open(good, 'rb').read())
    raw[8] = 2                                                # superblock version 2: new-style groups, not supported
    (tmp_path / 'v2.h5').write_bytes(raw)
    with pytest.raises(h5lite.H5Error, match='superblock version 2'):
        h5lite.File(str(tmp_path / 'v2.h5'))
    with pytest.raises(h5lite.H5Error):
        h5lite.File(good, 'w')
    with pytest.raises(h5lite.H5Error):
        h5lite.write(str(tmp_path / 'c.h5'), {'s': np.array(['a', 'b'])})
    truncated = tmp_path / 't.h5'
    raw[8] = 0
    truncated.write_bytes(bytes(raw[:len(raw) - 16]))
    with pytest.raises(h5lite.H5Error, match='past the end'):
        h5lite.File(str(truncated))['a'][:]


def _run_calls(loader):
    sys.path.insert(0, GOLD)
    from make_data_golden import CALLS, run_calls
    return CALLS, run_calls(loader)


def test_loader_matches_batches_recorded_from_the_reference_loader():
    gold = np.load(os.path.join(GOLD, 'data_batches.npz'))
    ld = AlexDataLoader(OPT)
    assert [ld.num_images, ld.num_channels, ld.seq_length, ld.vocab_size, len(ld.train_ix), len(ld.val_ix),
            len(ld.test_ix)] == gold['meta'].tolist()
    assert ld.getSeqLength() == 16 and ld.getVocabSize() == 50
    calls, out = _run_calls(ld)
    assert len(calls) == 22
    for k in gold.files:
        if k != 'meta':
            assert out[k].dtype == gold[k].dtype and np.array_equal(out[k], gold[k]), k
    ld.reset_iterator(0)
    assert ld.iterators[0] == 0
    with pytest.raises(ValueError):                           # my_utils.getopt: a required key is missing
        from imagecaptioning_b200.data import getopt
        getopt({'split': 0}, 'iterate')


def test_loader_live_against_reference_loader():
    """Build container only: the unmodified AlexCap/MyDataLoader.py (its `h5py.File` bound to h5lite.File, matplotlib
    absent -> empty module) and ours, same seeded call sequence, every returned tensor identical."""
    if not have_reference():
        pytest.skip('/root/reference not present on this machine')
    sys.path.insert(0, GOLD)
    from make_data_golden import reference_loader
    ref, ours = reference_loader(), AlexDataLoader(OPT)
    _, a = _run_calls(ref)
    _, b = _run_calls(ours)
    assert a.keys() == b.keys()
    for k in a:
        assert np.array_equal(a[k], b[k]), k
    for attr in AlexDataLoader.H5_KEYS:
        assert np.array_equal(getattr(ref, attr), getattr(ours, attr)), attr
    assert ref.idx_to_token == ours.idx_to_token and ref.train_ix == ours.train_ix and ref.test_ix == ours.test_ix


def test_feature_grid_source_and_wrap_rule():
    ld = AlexDataLoader(OPT, source='features')
    with h5lite.File(OPT['data_h5']) as f:
        feats, labels = f['features'][:], f['labels'][:]
    seen = []
    for _ in range(9):                                      # 30 train images, batch 4: 7 full batches, then the wrap
        x, y, info, _ = ld.get_batch({'split': 0, 'iterate': True}, 4)
        ix = [int(n[:5]) for n in info[0]['filename']]
        seen.append(ix)
        assert x.dtype == torch.float32 and np.array_equal(x.numpy(), feats[ix]) and np.array_equal(y.numpy(), labels[ix])
    assert seen[6] == [24, 25, 26, 27] and seen[7] == [28, 29] and seen[8] == [0, 1, 2, 3]   # MyDataLoader.py:69-73
    with pytest.raises(AssertionError):
        ld.get_batch({'split': 3}, 1)
    with pytest.raises(ValueError):
        AlexDataLoader(OPT, source='pixels')


def test_write_dataset_and_reference_checkpoint(tmp_path, reference_modules):
    from imagecaptioning_b200.modules import Decoder
    arrays = {k: np.zeros((3, 4), np.int32) for k in ('labels', 'attributes')}
    arrays.update({k: np.arange(3) for k in ('split', 'lengths', 'img_to_first_phr', 'img_to_last_phr')})
    arrays['split'] = np.array([0, 1, 2])
    arrays['features'] = np.ones((3, 2, 8), np.float32)
    info = {'idx_to_token': {'1': 'a', '2': 'b'}, 'idx_to_filename': {str(i): f'{i}.jpg' for i in range(3)},
            'attributes_labels': ['x'] * 4}
    write_dataset(str(tmp_path / 'd.h5'), str(tmp_path / 'd.json'), arrays, info)
    ld = AlexDataLoader({'data_h5': str(tmp_path / 'd.h5'), 'data_json': str(tmp_path / 'd.json'), 'device': 'cpu'},
                        source='features')
    assert (ld.num_images, ld.seq_length, ld.vocab_size, ld.val_ix) == (3, 4, 2, [1])
    # a checkpoint as train_LSTMwAttention.py saves it: AlexCapModel.state_dict() -> 'llm.*' + 'features.*' keys
    torch.manual_seed(123)
    ref = reference_modules['Decoder'](50, 16, 32, 24, 5, {}, 0.0, 'cpu')
    sd = {'llm.' + k: v for k, v in ref.state_dict().items()}
    sd['features.0.weight'] = torch.zeros(4, 3, 7, 7)
    torch.save(sd, str(tmp_path / 'ck.pth'))
    ours = Decoder(50, 16, 32, 24, 5, {}, 0.0, 'cpu')
    rest = load_reference_checkpoint(ours, str(tmp_path / 'ck.pth'))
    assert rest == ['features.0.weight']
    for k, v in ref.state_dict().items():
        assert torch.equal(ours.state_dict()[k], v), k
    assert load_reference_checkpoint(Decoder(50, 16, 32, 24, 5, {}, 0.0, 'cpu'), ref.state_dict()) == []


def test_h5lite_round_trip_property(tmp_path):
    """Random shapes / dtypes / byte orders / chunkings: write() -> File() returns the same arrays (hypothesis)."""
    from hypothesis import given, settings, strategies as st
    from hypothesis.extra import numpy as hnp

    dtypes = st.sampled_from(['<i1', '<u1', '<i2', '>i2', '<u2', '<i4', '>i4', '<u4', '<i8', '>u8', '<f4', '>f4', '<f8'])
    arrays = dtypes.flatmap(lambda dt: hnp.arrays(dt, hnp.array_shapes(min_dims=1, max_dims=4, min_side=0, max_side=6),
                                                  elements=None if 'f' not in dt else st.floats(-1e6, 1e6, width=32)))
    counter = [0]

    @settings(max_examples=40, deadline=None)
    @given(st.lists(arrays, min_size=0, max_size=9), st.sampled_from([0, 512]), st.booleans(), st.booleans())
    def run(arrs, userblock, deflate, shuffle):
        counter[0] += 1
        ds = {f'd{i}_{"x" * (i % 4)}': a for i, a in enumerate(arrs)}
        chunks = {k: tuple(max(1, s // 2 + 1) for s in a.shape) for k, a in ds.items() if a.size and a.ndim <= 3}
        p = str(tmp_path / f'p{counter[0]}.h5')
        h5lite.write(p, ds, userblock=userblock, chunks=chunks, deflate=deflate, shuffle=shuffle)
        with h5lite.File(p) as f:
            assert f.keys() == sorted(ds)
            for k, v in ds.items():
                a = f[k][:]
                assert a.shape == v.shape and a.dtype == v.dtype and np.array_equal(a, v), k

    run()
