"""CSV (+ optional wandb) logging with the reference's file layout (/root/reference/utils/logger.py:12-144):
`<save>/<env>/<experiment>/{train,val,eval}.csv` with the header fixed by the first row, `config.yaml` beside them,
resume = append. `ExperimentFromFile` in the reference later reads `step,success` from eval.csv."""
import os
from dataclasses import asdict, is_dataclass
from pathlib import Path

import yaml


class CsvLogger:
    def __init__(self, path, resume: bool = False):
        path = Path(path)
        path.parent.mkdir(parents=True, exist_ok=True)
        self.header = None
        if resume and path.exists() and path.stat().st_size > 0:
            with open(path, "r") as f:
                self.header = f.readline().strip().split(",")
        self.path = path
        self.file = open(path, "a" if resume else "w")

    def log(self, row: dict, step=None):
        if self.file is None:
            # a stopped experiment that trains again: SuccessiveHalving ranks every candidate ever scored, so a pruned
            # candidate can re-enter the population after Trainer has called stop() on it
            self.file = open(self.path, "a")
        row = dict(row)
        if step is not None:
            row["step"] = step
        if self.header is None:
            self.header = list(row)
            self.file.write(",".join(self.header) + "\n")
        self.file.write(",".join(str(row.get(k, "")) for k in self.header) + "\n")
        self.file.flush()

    def close(self):
        if self.file is not None:
            self.file.close()
            self.file = None


class Logger:
    def __init__(self, save_directory: Path, env_name: str, experiment_name: str, config, use_wandb: bool = False,
                 state_dict: dict | None = None):
        resume = state_dict is not None
        root = Path(save_directory) / env_name / experiment_name
        self.loggers = {g: CsvLogger(root / f"{g}.csv", resume=resume) for g in ("train", "val", "eval")}
        self.train_logger, self.val_logger, self.eval_logger = (self.loggers[g] for g in ("train", "val", "eval"))
        if not resume:
            cfg = asdict(config) if is_dataclass(config) else dict(config)
            with open(root / "config.yaml", "w") as f:
                yaml.dump(cfg, f)
        self.use_wandb = use_wandb
        self.wandb_id = state_dict.get("wandb_id") if resume else None
        self._wandb = None
        if use_wandb:
            import wandb  # optional dependency; only touched when asked for
            self._wandb = wandb
            run = wandb.init(project="fql", group=env_name, name=experiment_name, id=self.wandb_id,
                             resume="allow" if resume else None, config=asdict(config) if is_dataclass(config) else config)
            self.wandb_id = run.id

    def log(self, data: dict, step: int, group: str = "train"):
        self.loggers[group].log({k: (v.item() if hasattr(v, "item") else v) for k, v in data.items()}, step=step)
        if self._wandb is not None:
            self._wandb.log({f"{group}/{k}": v for k, v in data.items()}, step=step)

    def state_dict(self) -> dict:
        return {"wandb_id": self.wandb_id}

    def close(self):
        for lg in self.loggers.values():
            lg.close()
        if self._wandb is not None:
            self._wandb.finish()
            self._wandb = None
