"""Replay of a finished experiment from its `eval.csv` (reference: /root/reference/evaluator/experiment_from_file.py:9-54).

Same interface as `trainer.experiment.Experiment` as far as the strategy replay needs it: `train(n) -> finished`,
`evaluate() -> success at the current step`. The newest run directory `<config_to_name(cfg)>_<YYYYMMDD_HHMMSS>` under
`<save_directory>/<env_name>/` wins (reference :16-28). No pandas: the file is two columns of a CSV.
"""
from __future__ import annotations

import csv
import re
from pathlib import Path

from trainer.config import ExperimentConfig

_TS = re.compile(r"_(\d{8}_\d{6})")


def config_to_name(experiment_config: ExperimentConfig) -> str:
    """`field_value` of every tuned (non-None) field joined by `_` (reference :46-54)."""
    tuned = [(f, v) for f, v in vars(experiment_config).items() if v is not None]
    return "_".join(f"{f}_{v}" for f, v in tuned)


class ExperimentFromFile:
    def __init__(self, experiment_config: ExperimentConfig, save_directory: Path, env_name: str):
        self.save_directory = Path(save_directory)
        self.env_name = env_name

        def timestamp(p: Path) -> str:
            m = _TS.search(p.parent.name)
            return m.group(1) if m else ""

        files = list((self.save_directory / env_name).glob(f"{config_to_name(experiment_config)}_*/eval.csv"))
        if not files:
            # the reference fails with IndexError on sorted(...)[-1]; say what is missing instead
            raise FileNotFoundError(
                f"no eval.csv for {config_to_name(experiment_config)} under {self.save_directory / env_name}")
        path = sorted(files, key=timestamp)[-1]
        self.experiment_config = experiment_config
        self.current_step = 0
        self.steps, self.success = [], []
        with open(path, newline="") as f:
            for row in csv.DictReader(f):
                self.steps.append(int(float(row["step"])))
                self.success.append(float(row["success"]))
        self.max_steps = max(self.steps)

    def train(self, num_steps: int) -> bool:
        self.current_step += num_steps
        return self.current_step >= self.max_steps

    def evaluate(self, num_episodes: int = 50) -> float:
        hits = [s for st, s in zip(self.steps, self.success) if st == self.current_step]
        if len(hits) != 1:  # pandas .item() in the reference raises ValueError on 0 or >1 rows
            raise ValueError(f"{len(hits)} rows for step {self.current_step} in eval.csv")
        return hits[0]

    def get_label(self) -> str:
        return rf"$seed = {self.experiment_config.seed}, \alpha = {self.experiment_config.alpha:.2f}$"

    def get_data(self):
        """(steps, success) columns (the reference returns the same two columns as a DataFrame)."""
        return {"step": list(self.steps), "success": list(self.success)}
