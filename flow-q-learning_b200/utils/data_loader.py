"""Host data loaders with the reference's names and sampling semantics (/root/reference/utils/data_loader.py:8-47):
uniform rows (StepLoader), a random window of `sequence_length` steps inside a 1000-step episode (MultistepLoader),
and episode-start rows (InitialObservationLoader). They feed world-model training and the initial states of simulated
evaluation; the population update gathers on the GPU instead (fqlpop_set_dataset)."""
from abc import ABC, abstractmethod
from typing import Dict

import numpy as np

from fql.utils.datasets import Dataset

EPISODE_LENGTH = 1000


class DataLoader(ABC):
    @abstractmethod
    def sample(self, batch_size: int) -> Dict[str, np.ndarray]:
        ...


class StepLoader(DataLoader):
    def __init__(self, dataset: Dict[str, np.ndarray]):
        self.dataset = Dataset.create(**dataset)

    def sample(self, batch_size: int) -> Dict[str, np.ndarray]:
        return self.dataset.sample(batch_size)


class InitialObservationLoader(DataLoader):
    def __init__(self, dataset: Dict[str, np.ndarray]):
        starts = np.arange(0, len(dataset["observations"]), EPISODE_LENGTH)
        self.dataset = Dataset.create(**{k: v[starts] for k, v in dataset.items()})

    def sample(self, batch_size: int) -> Dict[str, np.ndarray]:
        return self.dataset.sample(batch_size)


class MultistepLoader(DataLoader):
    def __init__(self, dataset: Dict[str, np.ndarray], sequence_length: int = 128):
        self.sequence_length = sequence_length
        episodes = len(dataset["observations"]) // EPISODE_LENGTH
        self.dataset = Dataset.create(
            **{k: np.squeeze(np.reshape(v, (episodes, EPISODE_LENGTH, -1))) for k, v in dataset.items()})

    def sample(self, batch_size: int) -> Dict[str, np.ndarray]:
        eps = self.dataset.sample(batch_size)                                     # first RNG draw: episodes
        start = np.random.randint(0, EPISODE_LENGTH - self.sequence_length, size=batch_size)  # second: window starts
        t = start[:, None] + np.arange(self.sequence_length)[None, :]
        rows = np.arange(batch_size)[:, None]
        return {k: v[rows, t] for k, v in eps.items()}
