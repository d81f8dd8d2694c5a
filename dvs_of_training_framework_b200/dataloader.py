"""Loader factories -- signature-preserving mirror of the reference's `utils/dataloader.py`
(`choose_collate_function` :31-32, `get_common_dataset_params` :35-42, `get_trainset_params`
:45-59, `get_valset_params` :62-72, `get_dataset` :75-86, `get_dataloader` :89-108).  Pure wiring:
what changes is what gets wired in -- this package's `collate_wrapper` (optionally assembling the
batch directly in the 9 B/event storage form) and `PreprocessedDataloader` (pinned, undecoded
batches for `training.DevicePrefetcher`).  The map-style / iterable datasets that read the
per-frame HDF5 files (`utils/dataset.py:552-797`) stay the reference's: pass its `Dataset` /
`IterableDataset` classes as `dataset_classes`, or have `utils.dataset` importable."""
import functools
from types import SimpleNamespace

import torch

from .dataset import collate_wrapper
from .preprocessed import PreprocessedDataloader


def choose_collate_function(is_raw, encoded=False):
    """utils/dataloader.py:31-32; encoded=True: batches come out in the storage form (see collate_wrapper)."""
    if not is_raw:
        return None
    return functools.partial(collate_wrapper, encoded=True) if encoded else collate_wrapper


def get_common_dataset_params(args):
    return SimpleNamespace(shape=args.shape, batch_size=args.mbs, pin_memory=True, num_workers=args.num_workers,
                           min_seq_length=args.min_sequence_length, max_seq_length=args.max_sequence_length,
                           is_static_seq_length=not args.dynamic_sample_length)


def get_trainset_params(args, encoded=False):
    params = get_common_dataset_params(args)
    params.path = args.data_path / 'outdoor_day2'
    params.augmentation = True
    params.collapse_length = args.cl
    params.shuffle = True
    params.infinite = True
    params.is_raw = args.is_raw
    params.collate_fn = choose_collate_function(params.is_raw, encoded)
    params.preprocessed_dataset_path = args.preprocessed_dataset_path \
        if 'preprocessed_dataset_path' in vars(args) else None
    params.cache_dir = args.cache_dir if 'cache_dir' in vars(args) else None
    params.cache_size = args.cache_size
    return params


def get_valset_params(args, encoded=False):
    params = get_common_dataset_params(args)
    params.path = args.data_path / 'outdoor_day1'
    params.augmentation = False
    params.collapse_length = 1
    params.shuffle = False
    params.infinite = False
    params.is_raw = True
    params.collate_fn = choose_collate_function(params.is_raw, encoded)
    params.preprocessed_dataset_path = None
    return params


def _reference_dataset_classes():
    try:
        from utils.dataset import Dataset, IterableDataset           # the reference tree on sys.path
    except ImportError as e:
        raise RuntimeError('get_dataset needs the reference\'s HDF5 dataset classes: pass '
                           'dataset_classes=(Dataset, IterableDataset) or put the reference on sys.path') from e
    return Dataset, IterableDataset


def get_dataset(params, dataset_classes=None):
    kwargs = {'path': params.path, 'shape': params.shape, 'augmentation': params.augmentation,
              'collapse_length': params.collapse_length, 'is_raw': params.is_raw,
              'min_seq_length': params.min_seq_length, 'max_seq_length': params.max_seq_length,
              'is_static_seq_length': params.is_static_seq_length}
    Dataset, IterableDataset = dataset_classes or _reference_dataset_classes()
    if params.infinite:
        return IterableDataset(shuffle=params.shuffle, **kwargs)
    return Dataset(**kwargs)


def get_dataloader(params, sample_idx=0, process_only_once=True, dataset_classes=None, **preprocessed_kwargs):
    """utils/dataloader.py:89-108.  preprocessed_kwargs go to PreprocessedDataloader (device=, opener=)."""
    if params.preprocessed_dataset_path is not None:
        loader = PreprocessedDataloader(path=params.preprocessed_dataset_path, batch_size=params.batch_size,
                                        is_raw=params.is_raw, cache_dir=params.cache_dir,
                                        cache_size=params.cache_size, process_only_once=process_only_once,
                                        **preprocessed_kwargs)
        loader.set_index(sample_idx)
        return loader
    loader_kwargs = {}
    if not params.infinite:
        loader_kwargs['shuffle'] = params.shuffle
    return torch.utils.data.DataLoader(get_dataset(params, dataset_classes), collate_fn=params.collate_fn,
                                       batch_size=params.batch_size, num_workers=params.num_workers,
                                       pin_memory=params.pin_memory, **loader_kwargs)
