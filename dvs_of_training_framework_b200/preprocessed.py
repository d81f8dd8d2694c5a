"""The preprocessed-dataset loader of the reference (utils/dataset.py): index arithmetic over the
run-length encoded files (`select_batch_info_ranges` :19-65, `select_encoded_ranges` :67-115,
`select_quantized_ranges` :118-157), `join_batches` (:160-198), `read_data` / `read_encoded_batch` /
`read_encoded_quantized_batch` (:400-425, :505-549) and `PreprocessedDataloader` (:799-954) -- same
names, arguments and results, with two differences that are what this package is for:

  * what is read lands in PINNED host memory, one tensor per field, so the upload is asynchronous;
  * `PreprocessedDataloader.__next__` yields the batch in the STORAGE form it has on disk (9 B/event,
    uint8 frames) instead of expanding it to int64 on the host with Python loops
    (`decode_batch`, :335-373): hand it to `training.DevicePrefetcher`, which uploads it one step
    ahead and expands it on the device (`dataset.decode_batch`), or construct the loader with
    `device=...` to get the decoded training-form batch on that device directly (what the
    reference's loader returns, minus the host work).

HDF5 reading stays the reference's: files are opened with `h5py.File(name, 'r')` when h5py is
installed; any callable `opener(path) -> context manager yielding a nested mapping of sliceable
arrays` can be passed instead (the tests use .npz-backed stand-ins; h5py is absent offline)."""
import contextlib
from pathlib import Path

import numpy as np
import torch


def cumsum_with_prefix(tensor, dtype=None):
    """utils/common.py:26-40: [1, 2, 3] -> [0, 1, 3, 6]."""
    out = torch.zeros(tensor.numel() + 1, dtype=dtype or tensor.dtype)
    torch.cumsum(tensor.to(out.dtype), 0, out=out[1:])
    return out


def _span(begin, end):
    return {'begin': begin, 'end': end}


def select_batch_info_ranges(elements_per_sample, sample_begin, sample_end):
    """utils/dataset.py:19-65: [begin, end) of every per-batch tensor for samples [sample_begin, sample_end).
    A sample with e elements has e + 1 timestamps / images."""
    assert isinstance(sample_begin, int)
    assert isinstance(sample_end, int)
    assert sample_end > sample_begin
    shift = cumsum_with_prefix(elements_per_sample + 1, torch.long)
    t0, t1 = shift[sample_begin].item(), shift[sample_end].item()
    per_sample = ('idx', 'sequence_length', 'collapse_length', 'box', 'angle', 'is_flip')
    return {'timestamps': _span(t0, t1), 'elements_per_sample': _span(sample_begin, sample_end),
            'images': _span(t0, t1),
            'augmentation_params': {k: _span(sample_begin, sample_end) for k in per_sample}}


def select_encoded_ranges(events_per_element, elements_per_sample, sample_begin, sample_end):
    """utils/dataset.py:67-115: the same plus the event arrays (via the element and event prefix sums)."""
    assert isinstance(sample_begin, int)
    assert isinstance(sample_end, int)
    assert sample_end > sample_begin
    events_shift = cumsum_with_prefix(events_per_element, torch.long)
    elements_shift = cumsum_with_prefix(elements_per_sample, torch.long)
    el0, el1 = elements_shift[sample_begin].item(), elements_shift[sample_end].item()
    ev0, ev1 = events_shift[el0].item(), events_shift[el1].item()
    result = select_batch_info_ranges(elements_per_sample, sample_begin, sample_end)
    result['events'] = {k: _span(ev0, ev1) for k in ('x', 'y', 'timestamp', 'polarity')}
    result['events']['events_per_element'] = _span(el0, el1)
    return result


def select_quantized_ranges(channels_per_sample, elements_per_sample, sample_begin, sample_end):
    """utils/dataset.py:118-157: quantized files store `data` as (sum of channels, H, W)."""
    assert isinstance(sample_begin, int)
    assert isinstance(sample_end, int)
    assert sample_end > sample_begin
    shift = cumsum_with_prefix(channels_per_sample, torch.long)
    result = select_batch_info_ranges(elements_per_sample, sample_begin, sample_end)
    result['data'] = _span(shift[sample_begin].item(), shift[sample_end].item())
    result['channels_per_sample'] = _span(sample_begin, sample_end)
    return result


def _pin(t):
    return t.pin_memory() if torch.cuda.is_available() and t.numel() else t


def read_data(descriptor, ranges, pin=True):
    """utils/dataset.py:505-517: descriptor[k][begin:end] for every leaf of `ranges`, as tensors (pinned)."""
    def is_final(element):
        assert isinstance(element, dict), element
        return 'begin' in element and isinstance(element['begin'], int) and \
               'end' in element and isinstance(element['end'], int)
    assert isinstance(ranges, dict)
    result = {}
    for k, v in ranges.items():
        if is_final(v):
            t = torch.as_tensor(np.asarray(descriptor[k][v['begin']:v['end']]))
            result[k] = _pin(t) if pin else t.clone()
        else:
            result[k] = read_data(descriptor[k], v, pin)
    return result


def read_encoded_batch(descriptor, events_per_element, elements_per_sample, sample_begin, sample_end):
    """utils/dataset.py:400-425."""
    return read_data(descriptor, select_encoded_ranges(events_per_element, elements_per_sample,
                                                       sample_begin, sample_end))


def read_encoded_quantized_batch(descriptor, channels_per_sample, elements_per_sample, sample_begin, sample_end):
    """utils/dataset.py:520-549."""
    return read_data(descriptor, select_quantized_ranges(channels_per_sample, elements_per_sample,
                                                         sample_begin, sample_end))


def join_batches(batches):
    """utils/dataset.py:160-198: concatenate encoded batches field by field (pinned result)."""
    if len(batches) == 0:
        short = torch.short
        return {'events': {'x': torch.tensor([], dtype=short), 'y': torch.tensor([], dtype=short),
                           'timestamp': torch.tensor([], dtype=torch.float32),
                           'polarity': torch.tensor([], dtype=torch.bool),
                           'events_per_element': torch.tensor([], dtype=short)},
                'timestamps': torch.tensor([], dtype=torch.float32),
                'elements_per_sample': torch.tensor([], dtype=short),
                'images': torch.tensor([], dtype=torch.uint8), 'augmentation_params': {}}
    if len(batches) == 1:
        return batches[0]
    result = {}
    for k in batches[0].keys():
        if isinstance(batches[0][k], dict):
            result[k] = {sk: _pin(torch.cat([el[k][sk] for el in batches])) for sk in batches[0][k].keys()}
        elif batches[0][k] is None:
            assert k == 'augmentation_params'
            assert all([el[k] is None for el in batches])
            result[k] = None
        else:
            assert isinstance(batches[0][k], torch.Tensor)
            result[k] = _pin(torch.cat([el[k] for el in batches]))
    return result


def _h5_opener(path):
    try:
        import h5py
    except ImportError as e:                                   # no silent substitute for the file format
        raise RuntimeError('PreprocessedDataloader reads HDF5 through h5py (not installed); pass '
                           '`opener=` to read another container') from e
    return h5py.File(path, 'r')


class _FileCycle:
    """The part of utils/file_iterators.py the loader uses when no cache is configured: the files in
    order, again and again; `next()` -> an object with `.name` and `.release()`; `reset()`."""

    class _Handle:
        def __init__(self, name):
            self.name = name

        def release(self):
            pass

    def __init__(self, files):
        self.files, self.i = list(files), 0

    def next(self):
        h = self._Handle(self.files[self.i % len(self.files)])
        self.i += 1
        return h

    def reset(self):
        self.i = 0


class PreprocessedDataloader:
    """utils/dataset.py:799-954.  Iterates over the files of a preprocessed dataset (written by
    scripts/prepare_batches.py / quantize_preprocessed.py) and returns `batch_size` samples per
    `next()`, crossing file boundaries and wrapping around at the end.

    Attributes as in the reference: file_index / sample_index position, batch_size, files, length.
    `device=None` (default): the batch in the storage form, pinned -- feed `training.DevicePrefetcher`.
    `device='cuda'`: uploaded and decoded there (`dataset.decode_batch` / `decode_quantized_batch`):
    the training-form batch the reference's loader returns, as CUDA tensors."""

    def __init__(self, path, batch_size, is_raw, cache_dir=None, cache_size=0, process_only_once=True,
                 device=None, opener=None, suffix='.hdf5', file_iterator=None):
        path = Path(path)
        self.batch_size, self.is_raw, self.device = batch_size, is_raw, device
        self.opener = opener or _h5_opener
        self.files = sorted(path.glob('*' + suffix), key=lambda x: int(x.stem))
        assert len(self.files) > 0, f'No preprocessed dataset at {path} (on {suffix} files)'
        # cache_dir / cache_size / process_only_once configure the reference's background file copier
        # (utils/file_iterators.py, host I/O outside this package): pass its iterator as `file_iterator`
        self.iterator = file_iterator or _FileCycle(self.files)
        self.sample_index = 0
        self.length = sum(self._file2size(f, save_info=True) for f in self.files)
        self.current_file = self.iterator.next()

    def _open(self, filename):
        d = self.opener(filename)
        return d if hasattr(d, '__enter__') else contextlib.nullcontext(d)

    def _hdf5file2size(self, filename):
        with self._open(filename) as f:
            return len(f['elements_per_sample'])

    def _file2size(self, filename, save_info=False):
        """Samples in a file; cached in a `<stem>.info` YAML next to it like the reference."""
        import yaml
        filename = Path(filename)
        info = filename.parent / (filename.stem + '.info')
        if info.is_file():
            return yaml.safe_load(info.read_text())['size']
        size = self._hdf5file2size(filename)
        if save_info:
            try:
                info.write_text(yaml.dump({'size': size}))
            except OSError:
                pass                                                # read-only dataset directory
        return size

    def set_index(self, idx):
        """Moves the sample iterator to sample `idx` (modulo the dataset length)."""
        self.sample_index = idx % self.length
        self.current_file.release()
        self.iterator.reset()
        self.current_file = self.iterator.next()
        while True:
            file_size = self._file2size(self.current_file.name)
            if self.sample_index < file_size:
                break
            self.sample_index -= file_size
            self.current_file.release()
            self.current_file = self.iterator.next()

    def __len__(self):
        return self.length

    def __iter__(self):
        return self

    @staticmethod
    def _read_raw_batch(descriptor, begin, end):
        events_per_element = torch.as_tensor(np.asarray(descriptor['events']['events_per_element']))
        elements_per_sample = torch.as_tensor(np.asarray(descriptor['elements_per_sample']))
        return read_encoded_batch(descriptor, events_per_element, elements_per_sample, begin, end)

    @staticmethod
    def _read_quantized_batch(descriptor, begin, end):
        channels_per_sample = torch.as_tensor(np.asarray(descriptor['channels_per_sample']))
        elements_per_sample = torch.as_tensor(np.asarray(descriptor['elements_per_sample']))
        return read_encoded_quantized_batch(descriptor, channels_per_sample, elements_per_sample, begin, end)

    def __next__(self):
        num2read = self.batch_size
        batches = []
        read_fun = self._read_raw_batch if self.is_raw else self._read_quantized_batch
        while num2read > 0:
            left = self._file2size(self.current_file.name) - self.sample_index
            cur_num2read = min(left, num2read)
            next_sample_index = self.sample_index + cur_num2read
            if cur_num2read > 0:
                with self._open(self.current_file.name) as f:
                    batches.append(read_fun(f, self.sample_index, next_sample_index))
            self.sample_index = next_sample_index
            num2read -= cur_num2read
            if num2read > 0:
                self.current_file.release()
                self.current_file = self.iterator.next()
                self.sample_index = 0
        encoded_batch = join_batches(batches)
        if self.device is None:
            return encoded_batch
        from . import dataset

        def up(v):
            if isinstance(v, dict):
                return {k: up(x) for k, x in v.items()}
            return v.to(self.device, non_blocking=True) if isinstance(v, torch.Tensor) else v
        on_device = up(encoded_batch)
        return dataset.decode_batch(on_device) if self.is_raw else dataset.decode_quantized_batch(on_device)
