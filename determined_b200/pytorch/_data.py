"""``det.pytorch.DataLoader`` and the three batch-sampler adapters that make a data stream
resumable, shardable across data-parallel ranks and endless.

Parity: reference ``harness/determined/pytorch/_data.py:44-343``. Composition order is
Repeat → Distributed → Skip: the stream is first made endless, then every ``num_replicas``-th BATCH
(not sample) is dealt to each rank starting at ``rank``, then ``skip`` batches are dropped so a
restored trial continues exactly where its checkpoint left the stream.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, Iterator, List, Optional, Sequence, Union

import numpy as np
import torch
from torch.utils.data import BatchSampler, Dataset, RandomSampler, Sampler, SequentialSampler
from torch.utils.data import DataLoader as _TorchDataLoader


class RepeatBatchSampler(Sampler):
    """Endless: restarts the wrapped batch sampler (drawing a fresh shuffle) each time it runs out."""

    def __init__(self, batch_sampler: Sampler) -> None:
        self.batch_sampler = batch_sampler

    def __len__(self) -> int:
        return len(self.batch_sampler)  # type: ignore[arg-type]

    def __iter__(self) -> Iterator:
        while True:
            n = 0
            for batch in self.batch_sampler:
                n += 1
                yield batch
            if n == 0:
                raise RuntimeError("batch sampler produced no batches; cannot repeat an empty stream")


class DistributedBatchSampler(Sampler):
    """Deals whole batches round-robin: rank r gets batches r, r+n, r+2n, ..."""

    def __init__(self, batch_sampler: Sampler, num_workers: int, rank: int) -> None:
        if not 0 <= rank < num_workers:
            raise ValueError(f"rank {rank} must be in [0, {num_workers})")
        self.batch_sampler, self.num_workers, self.rank = batch_sampler, num_workers, rank

    def __len__(self) -> int:
        full = len(self.batch_sampler)  # type: ignore[arg-type]
        return full // self.num_workers + (1 if self.rank < full % self.num_workers else 0)

    def __iter__(self) -> Iterator:
        for i, batch in enumerate(self.batch_sampler):
            if i % self.num_workers == self.rank:
                yield batch


class SkipBatchSampler(Sampler):
    """Drops the first ``skip`` batches (fast-forward after a restore)."""

    def __init__(self, batch_sampler: Sampler, skip: int) -> None:
        self.batch_sampler, self.skip = batch_sampler, skip

    def __len__(self) -> int:
        return max(len(self.batch_sampler) - self.skip, 0)  # type: ignore[arg-type]

    def __iter__(self) -> Iterator:
        for i, batch in enumerate(self.batch_sampler):
            if i >= self.skip:
                yield batch


def adapt_batch_sampler(batch_sampler: Sampler, repeat: bool = False, skip: int = 0, num_replicas: int = 1,
                        rank: int = 0) -> Sampler:
    if repeat:
        batch_sampler = RepeatBatchSampler(batch_sampler)
    if num_replicas > 1:
        batch_sampler = DistributedBatchSampler(batch_sampler, num_replicas, rank)
    if skip > 0:
        batch_sampler = SkipBatchSampler(batch_sampler, skip)
    return batch_sampler


def _identity(batch: Any) -> Any:
    return batch


class _Prefetcher:
    """Runs the wrapped loader (batch gather, collation, page-locking) in a background thread, ``depth`` batches ahead
    of the training loop, preserving order. An extension over the reference (whose loaders rely on worker processes
    only): with CUDA graphs the training thread has little to do, so host-side batch assembly would otherwise be the
    serial bottleneck of large-image models."""

    def __init__(self, loader: Any, depth: int) -> None:
        self.loader, self.depth = loader, depth

    def __len__(self) -> int:
        return len(self.loader)

    def __iter__(self) -> Iterator:
        import queue
        import threading

        q: "queue.Queue" = queue.Queue(maxsize=self.depth)
        end = object()
        stop = threading.Event()

        def work() -> None:
            try:
                for item in self.loader:
                    while not stop.is_set():
                        try:
                            q.put(item, timeout=0.2)
                            break
                        except queue.Full:
                            continue
                    if stop.is_set():
                        return
                q.put(end)
            except BaseException as e:  # noqa: BLE001 - forwarded to the consumer
                q.put(e)

        t = threading.Thread(target=work, daemon=True, name="det-prefetch")
        t.start()
        _LIVE_PREFETCH.add((stop, t))
        try:
            while True:
                try:
                    item = q.get(timeout=1.0)
                except queue.Empty:
                    if stop.is_set() and not t.is_alive():
                        return                 # stopped from outside (interpreter exit): nothing more will arrive
                    continue
                if item is end:
                    return
                if isinstance(item, BaseException):
                    raise item
                yield item
        finally:
            # the thread must be OUT of torch before the interpreter goes away: a daemon thread killed inside a C++ frame
            # aborts the process ("terminate called without an active exception")
            stop.set()
            t.join(timeout=5)
            _LIVE_PREFETCH.discard((stop, t))


_LIVE_PREFETCH: set = set()


def stop_prefetch_threads() -> None:
    """Stop and join every live prefetch thread (iterators that were never closed: the training loader repeats forever).
    Registered to run at interpreter exit."""
    for stop, t in list(_LIVE_PREFETCH):
        stop.set()
    for stop, t in list(_LIVE_PREFETCH):
        t.join(timeout=5)
        _LIVE_PREFETCH.discard((stop, t))


import atexit as _atexit  # noqa: E402

_atexit.register(stop_prefetch_threads)


class DataLoader:
    """Holds the arguments of a ``torch.utils.data.DataLoader`` so the harness can build the real
    loader later with the repeat / shard / skip adapters. Map-style datasets only."""

    def __init__(self, dataset: Dataset, batch_size: Optional[int] = 1, shuffle: bool = False,
                 sampler: Optional[Sampler] = None, batch_sampler: Optional[Sampler] = None, num_workers: int = 0,
                 collate_fn: Optional[Callable] = None, pin_memory: bool = False, drop_last: bool = False,
                 timeout: float = 0, worker_init_fn: Optional[Callable] = None, multiprocessing_context: Any = None,
                 generator: Any = None, prefetch_batches: int = 0) -> None:
        if isinstance(dataset, torch.utils.data.IterableDataset):
            raise TypeError("det.pytorch.DataLoader supports map-style datasets only")
        if batch_sampler is not None:
            if batch_size != 1 or shuffle or sampler is not None or drop_last:
                raise ValueError("batch_sampler is mutually exclusive with batch_size, shuffle, sampler, drop_last")
        else:
            if sampler is not None and shuffle:
                raise ValueError("sampler is mutually exclusive with shuffle")
            if batch_size is None:
                raise ValueError("batch_size=None (no batching) is not supported")
            if sampler is None:
                sampler = RandomSampler(dataset, generator=generator) if shuffle else SequentialSampler(dataset)
            batch_sampler = BatchSampler(sampler, batch_size, drop_last)
        self.dataset = dataset
        self.batch_size = batch_size
        self.batch_sampler = batch_sampler
        self.num_workers = num_workers
        self.collate_fn = collate_fn
        self.pin_memory = pin_memory
        self.timeout = timeout
        self.worker_init_fn = worker_init_fn
        self.multiprocessing_context = multiprocessing_context
        self.prefetch_batches = int(prefetch_batches)

    def get_data_loader(self, repeat: bool = False, skip: int = 0, num_replicas: int = 1,
                        rank: int = 0) -> Any:
        bs = adapt_batch_sampler(self.batch_sampler, repeat=repeat, skip=skip, num_replicas=num_replicas, rank=rank)
        kwargs: Dict[str, Any] = dict(batch_sampler=bs, num_workers=self.num_workers, pin_memory=self.pin_memory,
                                      timeout=self.timeout, worker_init_fn=self.worker_init_fn)
        if self.collate_fn is not None:
            kwargs["collate_fn"] = self.collate_fn
        elif getattr(self.dataset, "det_batched", False) and hasattr(self.dataset, "__getitems__"):
            # the data set gathers a whole batch in one indexing operation and returns it already collated
            kwargs["collate_fn"] = _identity
        if self.multiprocessing_context is not None and self.num_workers > 0:
            kwargs["multiprocessing_context"] = self.multiprocessing_context
        loader = _TorchDataLoader(self.dataset, **kwargs)
        return _Prefetcher(loader, self.prefetch_batches) if self.prefetch_batches > 0 else loader

    def __iter__(self) -> Iterator:
        return iter(self.get_data_loader())

    def __len__(self) -> int:
        return len(self.batch_sampler)  # type: ignore[arg-type]


TorchData = Union[Dict[str, Any], Sequence[Any], torch.Tensor, np.ndarray]


def data_length(data: TorchData) -> int:
    """Number of records in a batch: length of the first tensor found."""
    if isinstance(data, (torch.Tensor, np.ndarray)):
        return int(data.shape[0])
    if isinstance(data, dict):
        if not data:
            raise ValueError("cannot infer the batch length of an empty dict")
        return data_length(next(iter(data.values())))
    if isinstance(data, (list, tuple)):
        if not data:
            raise ValueError("cannot infer the batch length of an empty sequence")
        return data_length(data[0])
    raise TypeError(f"cannot infer the batch length of {type(data).__name__}")


def to_device(data: Any, device: torch.device, warned_types: Optional[set] = None) -> Any:
    """Recursively move tensors / numpy arrays in dicts, lists and tuples to ``device`` (non-blocking, so
    pinned host batches overlap with compute)."""
    if isinstance(data, torch.Tensor):
        return data.to(device, non_blocking=True)
    if isinstance(data, np.ndarray):
        return torch.from_numpy(data).to(device, non_blocking=True)
    if isinstance(data, dict):
        return {k: to_device(v, device, warned_types) for k, v in data.items()}
    if isinstance(data, tuple) and hasattr(data, "_fields"):
        return type(data)(*(to_device(v, device, warned_types) for v in data))
    if isinstance(data, (list, tuple)):
        return type(data)(to_device(v, device, warned_types) for v in data)
    if warned_types is not None and type(data) not in warned_types:
        warned_types.add(type(data))
    return data
