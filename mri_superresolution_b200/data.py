"""Real-PNG input side of the drivers: the reference's ``MRISuperResDataset`` + ``DataLoader`` pair
(utils/dataset.py:13-136, scripts/train.py:193-233) restated for a path that consumes thousands of slices per second.

What is kept: pair discovery (sorted ``*.png`` of the full-resolution directory with a same-named low-resolution file,
utils/dataset.py:37-67; subject id from ``sub-<id>``, :52-58), ``random_split`` with ``torch.Generator().manual_seed(seed)``
(scripts/train.py:205-213: the SAME train / validation partition as the reference for a given seed), shuffled training order,
un-shuffled validation, ragged last batch (no ``drop_last``).
What changes: the host only DECODES.  A batch leaves the loader as two pinned uint8 tensors (low N x H x W, full N x 2H x 2W) --
a quarter of the float32 bytes over PCIe -- and ``ToTensor`` / ``augment_pair`` run on the device (utils/preprocess.py).  Decoding
uses a thread pool (PIL releases the GIL while inflating) and an optional in-memory cache of decoded slices (the reference's
per-item LRU ``cache_size``, :92-94, keyed the same way; here it holds the uint8 arrays, so a 100 k-slice set fits in host RAM).
"""
from __future__ import annotations

import os
import re
from concurrent.futures import ThreadPoolExecutor
from typing import Dict, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch

__all__ = ["PairedPNGDataset", "UInt8BatchLoader", "split_indices"]


class PairedPNGDataset:
    def __init__(self, full_res_dir: str, low_res_dir: str, cache: bool = True):
        self.full_res_dir, self.low_res_dir = str(full_res_dir), str(low_res_dir)
        # utils/dataset.py:37-39: every PNG of the full-resolution directory, sorted; __len__ counts these (:97-98)
        self.full_res_files = sorted(f for f in os.listdir(self.full_res_dir) if f.lower().endswith(".png"))
        self.valid_pairs, self.subjects, self.metadata = [], [], []
        for f in self.full_res_files:
            if os.path.exists(os.path.join(self.low_res_dir, f)):
                self.valid_pairs.append(f)
                m = re.search(r"sub-([A-Za-z0-9]+)", f)                               # :52
                self.subjects.append(m.group(1) if m else f)
                self.metadata.append({"filename": f, "subject": self.subjects[-1],
                                      "full_res_path": os.path.join(self.full_res_dir, f),
                                      "low_res_path": os.path.join(self.low_res_dir, f)})
        self._cache: Optional[Dict[int, Tuple[np.ndarray, np.ndarray]]] = {} if cache else None

    def __len__(self) -> int:
        return len(self.full_res_files)

    def load(self, idx: int) -> Tuple[np.ndarray, np.ndarray]:
        """(low uint8 H x W, full uint8 2H x 2W) of pair ``idx``: ``Image.open(...).convert('L')`` (:112-113)."""
        if self._cache is not None and idx in self._cache:
            return self._cache[idx]
        from PIL import Image
        name = self.full_res_files[idx]
        fp, lp = os.path.join(self.full_res_dir, name), os.path.join(self.low_res_dir, name)
        try:
            full = np.asarray(Image.open(fp).convert("L"))
            low = np.asarray(Image.open(lp).convert("L"))
        except Exception as exc:                                                       # :114-120: no silent placeholder
            raise RuntimeError(f"Error loading images for {name} at paths: {fp} and {lp}. Error: {exc}") from exc
        pair = (np.ascontiguousarray(low), np.ascontiguousarray(full))
        if self._cache is not None:
            self._cache[idx] = pair
        return pair

    def get_subject_indices(self, subject_id: str) -> List[int]:
        return [i for i, s in enumerate(self.subjects) if s == subject_id]             # :177-181


def split_indices(n: int, validation_split: float, seed: int) -> Tuple[List[int], List[int]]:
    """The partition ``random_split(dataset, [train, val], generator=torch.Generator().manual_seed(seed))`` makes
    (scripts/train.py:205-213): a seeded ``randperm`` cut at the train size."""
    val = int(validation_split * n)
    perm = torch.randperm(n, generator=torch.Generator().manual_seed(seed)).tolist()
    return perm[: n - val], perm[n - val:]


class UInt8BatchLoader:
    """Iterable of ``(low_u8 N x H x W, full_u8 N x 2H x 2W)`` pinned CPU tensors over ``indices`` of a PairedPNGDataset.

    ``rank`` / ``world`` (data parallel): every global batch of ``batch_size * world`` pairs is cut into ``world`` equal shards
    and rank r takes the r-th; an incomplete last global batch is dropped so that all ranks see the same number of equally shaped
    batches (drivers.train requires it).  Single-process (world = 1) keeps the ragged last batch like the reference's DataLoader."""

    def __init__(self, dataset: PairedPNGDataset, indices: Sequence[int], batch_size: int, shuffle: bool, seed: int = 0,
                 num_workers: int = 8, rank: int = 0, world: int = 1, pin: bool = True):
        self.ds, self.indices, self.bs, self.shuffle = dataset, list(indices), int(batch_size), shuffle
        self.rank, self.world, self.pin = rank, world, pin and torch.cuda.is_available()
        self._gen = torch.Generator().manual_seed(seed)
        self._pool = ThreadPoolExecutor(max_workers=max(1, num_workers)) if num_workers > 0 else None

    def __len__(self) -> int:
        n = len(self.indices)
        if self.world > 1:
            return n // (self.bs * self.world)
        return (n + self.bs - 1) // self.bs

    def _batch(self, idxs: Sequence[int]):
        pairs = list(self._pool.map(self.ds.load, idxs)) if self._pool is not None else [self.ds.load(i) for i in idxs]
        shapes = {(p[0].shape, p[1].shape) for p in pairs}
        if len(shapes) != 1:
            raise RuntimeError(f"slices of one batch differ in size: {sorted(shapes)} (the reference's default collate fails here too)")
        low = torch.from_numpy(np.stack([p[0] for p in pairs]))
        full = torch.from_numpy(np.stack([p[1] for p in pairs]))
        if self.pin:
            low, full = low.pin_memory(), full.pin_memory()
        return low, full

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        order = self.indices
        if self.shuffle:
            perm = torch.randperm(len(order), generator=self._gen).tolist()
            order = [order[i] for i in perm]
        if self.world > 1:
            g = self.bs * self.world
            for s in range(0, (len(order) // g) * g, g):
                yield self._batch(order[s + self.rank * self.bs: s + (self.rank + 1) * self.bs])
        else:
            for s in range(0, len(order), self.bs):
                yield self._batch(order[s: s + self.bs])
