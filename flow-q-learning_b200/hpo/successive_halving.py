"""Successive halving over the alpha population. Semantics follow the reference bit for bit
(/root/reference/hpo/successive_halving.py:11-115; known-answer vectors in tests/golden/successive_halving.json
were produced by executing the reference file):

  * milestones: eta = 1/(1-f); T = int((total - h)(1-f)); R = int(floor(ln N / ln eta)); int(T eta^-r) + h, r < R
  * update appends the score; performed_evaluations = longest history seen
  * sample prunes only when performed_evaluations is a milestone, >= h and more than one candidate is left; it
    keeps max(1, int(n - n f)) candidates ranked by sum(last h scores) / h -- always divided by h, and ranked over
    EVERY candidate ever scored (eliminated ones keep their frozen history and may re-enter). Ties keep the
    insertion order of candidate_scores (stable sort).
"""
from collections import defaultdict
from typing import List, Optional, Set

import numpy as np

from hpo.strategy import HpoStrategy
from trainer.config import ExperimentConfig


class SuccessiveHalving(HpoStrategy):
    def __init__(self, population: Set[ExperimentConfig], total_evaluations: int, fraction: float = 0.5,
                 history_length: int = 1, state_dict: Optional[dict] = None) -> None:
        super().__init__(population, total_evaluations, state_dict)
        self.fraction = fraction
        self.history_length = history_length
        self.candidate_scores = defaultdict(list)
        self.performed_evaluations = 0
        self.halving_milestones = self.compute_halving_milestones()
        if state_dict is not None:
            self.candidate_scores = state_dict["candidate_scores"]
            self.performed_evaluations = state_dict["performed_evaluations"]

    def state_dict(self) -> dict:
        out = super().state_dict()
        out["candidate_scores"] = dict(self.candidate_scores)
        out["performed_evaluations"] = self.performed_evaluations
        return out

    def update(self, candidate: ExperimentConfig, performance: float) -> None:
        if candidate not in self.candidate_scores:
            self.candidate_scores[candidate] = []
        history = self.candidate_scores[candidate]
        history.append(performance)
        if len(history) > self.performed_evaluations:
            self.performed_evaluations = len(history)

    def sample(self) -> Set[ExperimentConfig]:
        at_milestone = self.performed_evaluations in self.halving_milestones
        if not at_milestone or self.performed_evaluations < self.history_length or len(self.population) <= 1:
            return self.population
        n = len(self.population)
        keep = max(1, int(n - n * self.fraction))
        h = self.history_length
        ranked = sorted(((sum(scores[-h:]) / h, cand) for cand, scores in self.candidate_scores.items()),
                        key=lambda item: item[0], reverse=True)
        self.population = {cand for _, cand in ranked[:keep]}
        return self.population

    def compute_halving_milestones(self) -> List[int]:
        n = len(self.population)
        keep_frac = 1 - self.fraction
        eta = 1 / keep_frac
        budget = int((self.total_evaluations - self.history_length) * keep_frac)
        rounds = int(np.floor(np.log(n) / np.log(eta)))  # numpy's log, as the reference: the quotient decides R
        return [int(budget * eta ** -r) + self.history_length for r in range(rounds)]
