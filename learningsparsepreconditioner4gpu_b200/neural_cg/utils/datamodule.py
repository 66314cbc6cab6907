"""Host-side mirror of ``neural_cg/utils/datamodule.py``: ``FolderDataModule`` with the reference's 80/20 split
(``train_test_split(random_state=42, shuffle=True)``, :25-27) and loader names (:47-60).  Lightning and PyG's
``DataLoader`` are absent: a loader here is a plain re-iterable that yields ``Data`` batches built with
``neural_cg.data.collate`` (block-diagonal batching, ``ptr`` / ``batch`` as PyG makes them)."""
from __future__ import annotations

from typing import List, Sequence

import numpy as np
from sklearn.model_selection import train_test_split

from ..data import collate
from ..dataset import FolderDataset


def split_dataset(dataset_len: int, train_ratio: float):
    train_idx, val_idx = train_test_split(range(dataset_len), train_size=train_ratio, random_state=42, shuffle=True)
    return list(train_idx), list(val_idx)


class Loader:
    """Re-iterable over ``indices`` of ``dataset`` in batches (``shuffle`` draws a fresh permutation per epoch)."""

    def __init__(self, dataset, indices: Sequence[int], batch_size: int = 1, shuffle: bool = False, seed: int = 0):
        self.dataset, self.indices, self.batch_size, self.shuffle = dataset, list(indices), int(batch_size), shuffle
        self._rng = np.random.default_rng(seed)

    def __len__(self) -> int:
        return (len(self.indices) + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        order = self._rng.permutation(len(self.indices)).tolist() if self.shuffle else range(len(self.indices))
        batch: List = []
        for k in order:
            batch.append(self.dataset.get(self.indices[k]))
            if len(batch) == self.batch_size:
                yield collate(batch)
                batch = []
        if batch:
            yield collate(batch)


class FolderDataModule:
    def __init__(self, data_config, split_config, batch_size):
        cfg = data_config.to_container() if hasattr(data_config, "to_container") else dict(data_config)
        self.dataset = FolderDataset(**cfg)
        self.batch_size = int(batch_size)
        train_ratio = split_config["train"]
        self.train_idx, self.val_idx = split_dataset(len(self.dataset), train_ratio)

    def inference_dataloader(self):
        return Loader(self.dataset, range(len(self.dataset)), 1, False)   # the entire dataset

    def train_dataloader(self):
        return Loader(self.dataset, self.train_idx, self.batch_size, True)

    def val_dataloader(self):
        return Loader(self.dataset, self.val_idx, 1, False)

    def test_dataloader(self):
        return self.val_dataloader()
