"""Task interface (/root/reference/task/task.py:5-45): offline batches for training + an environment to evaluate in."""
from abc import ABC, abstractmethod


class Task(ABC):
    @abstractmethod
    def sample(self, dataset, batch_size: int):
        """Batch dict from the 'train' or 'val' dataset."""

    @abstractmethod
    def reset(self, seed: int = 0):
        """-> (observations [n, O], infos)"""

    @abstractmethod
    def step(self, action):
        """-> (next_observations, rewards, terminated, truncated, infos)"""

    @abstractmethod
    def close(self):
        """Release the evaluation environments."""
