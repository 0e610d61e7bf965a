"""Population-pruning strategy interface (/root/reference/hpo/strategy.py:7-58)."""
from abc import ABC, abstractmethod
from typing import Optional, Set

from trainer.config import ExperimentConfig


class HpoStrategy(ABC):
    def __init__(self, population: Set[ExperimentConfig], total_evaluations: int, state_dict: Optional[dict] = None):
        self.population = population if state_dict is None else state_dict["population"]
        self.total_evaluations = total_evaluations

    def state_dict(self) -> dict:
        return {"population": self.population}

    @abstractmethod
    def update(self, candidate: ExperimentConfig, performance: float) -> None:
        """Record one evaluation score of `candidate`."""

    @abstractmethod
    def sample(self) -> Set[ExperimentConfig]:
        """Candidates that keep training."""
