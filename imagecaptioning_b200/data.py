"""Input side of the hot path: the reference's on-disk formats.

`AlexDataLoader` mirrors `AlexCap/MyDataLoader.py:9-95` member for member -- same constructor `opt` keys (`data_h5`,
`data_json`, `device`), same attributes (`info`, `vocab_size`, `idx_to_token`, `labels`, `lengths`, `split`, `attributes`,
`img_to_first_phr`, `img_to_last_phr`, `num_images`, `num_channels`, `seq_length`, `iterators`, `train_ix/val_ix/test_ix`),
same `get_batch(opt, batch_size, idx=-1)` return tuple and iterator arithmetic (including the wrap-to-0 rule of `:69-72`
and the `torch.multinomial` draw of `:76`, so a seeded run visits the same images) -- but reads the HDF5 file with
`h5lite` (h5py / libhdf5 are not in this image) and can hand out precomputed CNN feature grids instead of pixels:

* `source='images'` (the reference's behaviour): uint8 NHWC -> NCHW -> torchvision ResNet-101 preset transform.
* `source='features'`: dataset `/features` `[M, L, D]` float (north_star's "precomputed CNN feature grids"), returned as-is,
  optionally as pinned host tensors (`pin_memory=True`) so `DecoderTrainStep.prefetch()` can overlap the H2D copy.

Checkpoints: `load_reference_checkpoint()` reads what `train_LSTM.py:123` / `train_LSTMwAttention.py` save
(`torch.save(model.state_dict())` of an `AlexCapModel`: decoder keys are prefixed `llm.`, the CNN's `features.`) and
loads the decoder part into this package's `Decoder` / `LanguageModel`.
"""
import json

import numpy as np
import torch

from . import h5lite


def getopt(opt, key, default_value=None):
    """AlexCap/my_utils.py:5-8."""
    if default_value is None and (opt is None or key not in opt):
        raise ValueError(f'Error: required key {key} was not provided in an opt.')
    return opt[key] if opt and key in opt else default_value


def _resnet_preprocess():
    import torchvision                                               # MyDataLoader.py:38
    return torchvision.models.ResNet101_Weights.IMAGENET1K_V2.transforms()


class AlexDataLoader:
    H5_KEYS = ['img_to_first_phr', 'img_to_last_phr', 'labels', 'lengths', 'split', 'attributes']   # MyDataLoader.py:26

    def __init__(self, opt, source='images', pin_memory=False, preprocess=None):
        if source not in ('images', 'features'):
            raise ValueError("source must be 'images' or 'features'")
        self.h5_file_path = getopt(opt, 'data_h5', 'AlexCap/data/face2text-data.h5')
        self.json_file_path = getopt(opt, 'data_json', 'AlexCap/data/face2text-dicts.json')
        self.device = opt['device']
        self.source, self.pin_memory = source, pin_memory
        with open(self.json_file_path, 'r') as f:
            self.info = json.load(f)
        self.vocab_size = len(self.info['idx_to_token'])
        self.idx_to_token = list(self.info['idx_to_token'].items())
        self.h5_file = h5lite.File(self.h5_file_path, 'r')
        for key in self.H5_KEYS:
            setattr(self, key, self.h5_file['/' + key][:])
        if source == 'images' or 'images' in self.h5_file:
            images_size = self.h5_file['images'].shape
            assert len(images_size) == 4, '/images should be a 4D tensor'
            self.num_images, self.num_channels = images_size[0], images_size[3]
        else:
            self.num_images, self.num_channels = self.h5_file['features'].shape[0], 0
        if source == 'features':
            fs = self.h5_file['features'].shape
            assert len(fs) == 3 and fs[0] == self.num_images, '/features should be [num_images, L, D]'
        self.seq_length = self.h5_file['labels'].shape[1]
        self.iterators = {0: 0, 1: 0, 2: 0}
        self.attributes_labels = self.info['attributes_labels']
        self.preprocess = preprocess if preprocess is not None else (_resnet_preprocess() if source == 'images' else None)
        split = np.asarray(self.split)
        self.train_ix = np.nonzero(split == 0)[0].tolist()            # :40-46, without one HDF5 read per image
        self.val_ix = np.nonzero(split == 1)[0].tolist()
        self.test_ix = np.nonzero(split == 2)[0].tolist()

    def getSeqLength(self):
        return self.seq_length

    def getVocabSize(self):
        return self.vocab_size

    def reset_iterator(self, split_val):
        self.iterators[split_val] = 0

    def _place(self, t):
        if self.pin_memory:
            return t.pin_memory()
        return t.to(self.device)

    def get_batch(self, opt, batch_size, idx=-1):
        split_val = getopt(opt, 'split', 0)
        iterate = getopt(opt, 'iterate', True)
        assert split_val in (0, 1, 2), 'split must be integer, either 0 (train), 1 (val) or 2 (test)'
        split_ix = (self.train_ix, self.val_ix, self.test_ix)[split_val]
        assert len(split_ix) > 0, 'split is empty ?'
        max_index = len(split_ix)
        if iterate:                                                   # :67-73
            ri = self.iterators[split_val]
            ri_next = ri + batch_size
            if ri_next >= max_index:
                ri_next = 0
            self.iterators[split_val] = ri_next
            ix = split_ix[ri:ri + batch_size]
        elif idx == -1:                                               # :75-77
            ri = np.sort(torch.multinomial(torch.ones(max_index), batch_size).cpu().numpy())
            ix = [split_ix[r] for r in ri]
        else:                                                         # :78-80
            ix = [split_ix[idx]]
            ri = idx
        if self.source == 'images':
            img = torch.from_numpy(self.h5_file['/images'][ix, :, :, :self.num_channels]).permute(0, 3, 1, 2)
            img = self.preprocess(img)
        else:
            img = torch.from_numpy(np.ascontiguousarray(self.h5_file['/features'][ix])).float()
            if self.preprocess is not None:
                img = self.preprocess(img)
        label_array = torch.from_numpy(self.labels[ix])
        attr_array = torch.clamp(torch.from_numpy(self.attributes[ix]), 0)
        filename = [self.info['idx_to_filename'][str(i)] for i in ix]
        info_table = [{'filename': filename, 'split_bounds': [ri, max_index]}]
        return self._place(img), self._place(label_array), info_table, attr_array


def write_dataset(h5_path, json_path, arrays, info):
    """Write the pair of files `my_model_preprocess.py:296-329` produces (HDF5 datasets + the dictionaries JSON)."""
    h5lite.write(h5_path, arrays)
    with open(json_path, 'w') as f:
        json.dump(info, f)


def load_reference_checkpoint(llm, path_or_state, strict=True):
    """Load a reference checkpoint (`torch.save(model.state_dict(), ...)`, train_LSTM.py:123 / :34) into `llm`.
    Accepts an `AlexCapModel` state dict (keys `llm.*`, `features.*`, `crit.*`) or a bare decoder state dict.
    Returns the keys that were not consumed (the CNN's)."""
    sd = path_or_state
    if not isinstance(sd, dict):
        sd = torch.load(path_or_state, map_location='cpu', weights_only=True)
    own = {k[len('llm.'):]: v for k, v in sd.items() if k.startswith('llm.')}
    rest = [k for k in sd if not k.startswith('llm.')]
    if not own:                                                       # a bare decoder state dict
        own, rest = dict(sd), []
    llm.load_state_dict(own, strict=strict)
    return rest
