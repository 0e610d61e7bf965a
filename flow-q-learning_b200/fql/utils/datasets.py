"""Offline dataset containers with the call surface the reference uses from the (absent) fql submodule:
Dataset.create(**arrays), .sample(batch_size), .size, dict(dataset), ReplayBuffer.create_from_initial_dataset(dict, size)
(/root/reference/task/offline_task_simulated.py:34-58, /root/reference/utils/data_loader.py:19,31,44).

Host sampling is uniform with replacement from numpy's global RNG -- `np.random.randint(size, size=B)` then a fancy-index
gather per array [UPSTREAM-RECALL seohongpark/fql utils/datasets.py] -- which keeps `np.random.seed` / the Trainer's
checkpointed RNG state meaningful. The population hot path does NOT go through here: it binds the same arrays in HBM
(`device_arrays`) and gathers on the GPU (csrc/pointwise.cu prep_kernel).
"""
from __future__ import annotations

from typing import Dict

import numpy as np


class Dataset(dict):
    """Dict of equally long row-major arrays."""

    @classmethod
    def create(cls, freeze: bool = True, **fields) -> "Dataset":
        data = {k: np.asarray(v) for k, v in fields.items()}
        n = {len(v) for v in data.values()}
        if len(n) != 1:
            raise ValueError(f"dataset arrays disagree on length: { {k: len(v) for k, v in data.items()} }")
        if freeze:
            for v in data.values():
                v.setflags(write=False)
        ds = cls(data)
        ds.size = n.pop()
        return ds

    def get_random_idxs(self, num_idxs: int) -> np.ndarray:
        return np.random.randint(self.size, size=num_idxs)

    def get_subset(self, idxs) -> Dict[str, np.ndarray]:
        return {k: v[idxs] for k, v in self.items()}

    def sample(self, batch_size: int, idxs=None) -> Dict[str, np.ndarray]:
        if idxs is None:
            idxs = self.get_random_idxs(batch_size)
        return self.get_subset(idxs)

    def device_arrays(self, device, keys=("observations", "actions", "next_observations", "rewards", "masks")):
        """fp32 copies of the first `size` rows in device memory (what fqlpop_set_dataset binds)."""
        import torch
        return [torch.as_tensor(np.ascontiguousarray(self[k][: self.size], dtype=np.float32)).to(device) for k in keys]


class ReplayBuffer(Dataset):
    """Fixed-capacity buffer pre-filled with an offline dataset; `size` counts the valid rows."""

    @classmethod
    def create_from_initial_dataset(cls, init_dataset: dict, size: int) -> "ReplayBuffer":
        n = len(next(iter(init_dataset.values())))
        if size < n:
            raise ValueError("buffer smaller than the initial dataset")
        data = {}
        for k, v in init_dataset.items():
            v = np.asarray(v)
            buf = np.zeros((size, *v.shape[1:]), dtype=v.dtype)
            buf[:n] = v
            data[k] = buf
        rb = cls(data)
        rb.max_size = size
        rb.size = rb.pointer = n
        return rb

    def add_transition(self, transition: dict) -> None:
        for k, v in transition.items():
            self[k][self.pointer] = v
        self.pointer = (self.pointer + 1) % self.max_size
        self.size = max(self.pointer, self.size)

    def clear(self) -> None:
        self.size = self.pointer = 0
