"""paddle.io (reference: python/paddle/io/{reader,dataloader/*}.py).  torch.utils.data does worker processes and
pinned-memory staging; this module restores Paddle's signatures, its BatchSampler/DistributedBatchSampler
semantics (padding to a multiple of nranks, ``set_epoch`` reshuffle; dataloader/batch_sampler.py:187-380) and
list-returning batches placed on the current device."""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.utils.data as tud

from ..framework.place import current_torch_device

Dataset = tud.Dataset
IterableDataset = tud.IterableDataset
Subset = tud.Subset
ConcatDataset = tud.ConcatDataset
ChainDataset = tud.ChainDataset
get_worker_info = tud.get_worker_info
Sampler = tud.Sampler


class TensorDataset(Dataset):
    def __init__(self, tensors):
        self.tensors = tensors

    def __getitem__(self, i):
        return tuple(t[i] for t in self.tensors)

    def __len__(self):
        return self.tensors[0].shape[0]


class ComposeDataset(Dataset):
    def __init__(self, datasets):
        self.datasets = list(datasets)

    def __len__(self):
        return len(self.datasets[0])

    def __getitem__(self, i):
        out = []
        for d in self.datasets:
            s = d[i]
            out.extend(s if isinstance(s, (tuple, list)) else [s])
        return tuple(out)


def random_split(dataset, lengths, generator=None):
    return tud.random_split(dataset, lengths, generator=generator)


class SequenceSampler(Sampler):
    def __init__(self, data_source):
        self.data_source = data_source

    def __iter__(self):
        return iter(range(len(self.data_source)))

    def __len__(self):
        return len(self.data_source)


class RandomSampler(Sampler):
    def __init__(self, data_source, replacement=False, num_samples=None, generator=None):
        self.data_source, self.replacement, self._num, self.generator = data_source, replacement, num_samples, generator

    @property
    def num_samples(self):
        return self._num if self._num is not None else len(self.data_source)

    def __iter__(self):
        n = len(self.data_source)
        if self.replacement:
            return iter(np.random.randint(0, n, self.num_samples).tolist())
        return iter(np.random.permutation(n)[: self.num_samples].tolist())

    def __len__(self):
        return self.num_samples


class SubsetRandomSampler(Sampler):
    def __init__(self, indices, generator=None):
        self.indices, self.generator = list(indices), generator

    def __iter__(self):
        import torch as _t
        for i in _t.randperm(len(self.indices), generator=self.generator).tolist():
            yield self.indices[i]

    def __len__(self):
        return len(self.indices)


class WeightedRandomSampler(Sampler):
    def __init__(self, weights, num_samples, replacement=True):
        self.weights, self.num_samples, self.replacement = np.asarray(weights, dtype=np.float64), num_samples, replacement

    def __iter__(self):
        p = self.weights / self.weights.sum()
        return iter(np.random.choice(len(p), self.num_samples, replace=self.replacement, p=p).tolist())

    def __len__(self):
        return self.num_samples


class BatchSampler(Sampler):
    def __init__(self, dataset=None, sampler=None, shuffle=False, batch_size=1, drop_last=False):
        if dataset is None and sampler is None:
            raise ValueError("either dataset or sampler should be set")
        self.sampler = sampler if sampler is not None else (RandomSampler(dataset) if shuffle else SequenceSampler(dataset))
        self.batch_size, self.drop_last = batch_size, drop_last

    def __iter__(self):
        batch = []
        for i in self.sampler:
            batch.append(i)
            if len(batch) == self.batch_size:
                yield batch
                batch = []
        if batch and not self.drop_last:
            yield batch

    def __len__(self):
        n = len(self.sampler)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size


class DistributedBatchSampler(BatchSampler):
    def __init__(self, dataset, batch_size, num_replicas=None, rank=None, shuffle=False, drop_last=False):
        from ..distributed import get_rank, get_world_size
        self.dataset, self.batch_size, self.shuffle, self.drop_last = dataset, batch_size, shuffle, drop_last
        self.nranks = num_replicas if num_replicas is not None else get_world_size()
        self.local_rank = rank if rank is not None else get_rank()
        self.epoch = 0
        self.num_samples = int(math.ceil(len(dataset) * 1.0 / self.nranks))
        self.total_size = self.num_samples * self.nranks

    def __iter__(self):
        n = len(self.dataset)
        indices = np.arange(n).tolist()
        indices += indices[: (self.total_size - len(indices))]
        if self.shuffle:
            np.random.RandomState(self.epoch).shuffle(indices)
            self.epoch += 1

        def subsample(ind):
            out = []
            last = self.total_size % (self.batch_size * self.nranks)
            for i in range(self.local_rank * self.batch_size, len(ind) - last, self.batch_size * self.nranks):
                out.extend(ind[i:i + self.batch_size])
            ind = ind[len(ind) - last:]
            per = last // self.nranks
            out.extend(ind[self.local_rank * per:(self.local_rank + 1) * per])
            return out

        if self.nranks > 1:
            indices = subsample(indices)
        batch = []
        for i in indices:
            batch.append(i)
            if len(batch) == self.batch_size:
                yield batch
                batch = []
        if batch and not self.drop_last:
            yield batch

    def __len__(self):
        n = self.num_samples
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def set_epoch(self, epoch):
        self.epoch = epoch


def default_collate_fn(batch):
    s = batch[0]
    if isinstance(s, torch.Tensor):
        return torch.stack(batch, 0)
    if isinstance(s, np.ndarray):
        return torch.from_numpy(np.stack(batch, 0))
    if isinstance(s, (int, np.integer)):
        return torch.tensor(batch, dtype=torch.int64)
    if isinstance(s, (float, np.floating)):
        return torch.tensor(batch, dtype=torch.float32)
    if isinstance(s, (str, bytes)):
        return batch
    if isinstance(s, dict):
        return {k: default_collate_fn([b[k] for b in batch]) for k in s}
    if isinstance(s, (list, tuple)):
        return [default_collate_fn(list(x)) for x in zip(*batch)]
    raise TypeError(f"batch data can only contains: tensor, numpy.ndarray, dict, list, number, but got {type(s)}")


def default_convert_fn(sample):
    """``batch_size=None``: no batching, the fields of one sample become tensors as they are (reference:
    python/paddle/io/dataloader/collate.py default_convert_fn)."""
    if isinstance(sample, torch.Tensor):
        return sample
    if isinstance(sample, np.ndarray):
        return torch.from_numpy(np.array(sample))
    if isinstance(sample, (str, bytes)):
        return sample
    if isinstance(sample, dict):
        return {k: default_convert_fn(v) for k, v in sample.items()}
    if isinstance(sample, (list, tuple)):
        return [default_convert_fn(v) for v in sample]
    return sample


class DataLoader:
    def __init__(self, dataset, feed_list=None, places=None, return_list=True, batch_sampler=None, batch_size=1,
                 shuffle=False, drop_last=False, collate_fn=None, num_workers=0, use_buffer_reader=True,
                 prefetch_factor=2, use_shared_memory=True, timeout=0, worker_init_fn=None, persistent_workers=False):
        self.dataset, self.batch_sampler = dataset, batch_sampler
        self.return_list = return_list
        self._device = current_torch_device()
        kw = dict(num_workers=num_workers, collate_fn=collate_fn or default_collate_fn, timeout=timeout,
                  worker_init_fn=worker_init_fn, pin_memory=self._device.type == "cuda")
        if num_workers > 0:
            kw.update(prefetch_factor=prefetch_factor, persistent_workers=persistent_workers)
        if isinstance(dataset, tud.IterableDataset):
            self._loader = tud.DataLoader(dataset, batch_size=batch_size, drop_last=drop_last, **kw)
        else:
            if batch_sampler is None:
                if batch_size is None:
                    kw["collate_fn"] = collate_fn or default_convert_fn
                    self._loader = tud.DataLoader(dataset, batch_size=None, shuffle=shuffle, **kw)
                    return
                batch_sampler = BatchSampler(dataset, shuffle=shuffle, batch_size=batch_size, drop_last=drop_last)
                self.batch_sampler = batch_sampler
            self._loader = tud.DataLoader(dataset, batch_sampler=batch_sampler, **kw)

    def _place(self, x):
        if isinstance(x, torch.Tensor):
            return x.to(self._device, non_blocking=True)
        if isinstance(x, dict):
            return {k: self._place(v) for k, v in x.items()}
        if isinstance(x, (list, tuple)):
            return [self._place(v) for v in x]
        return x

    def __iter__(self):
        for batch in self._loader:
            yield self._place(batch)

    def __len__(self):
        return len(self._loader)

    def __call__(self):
        return self.__iter__()
