"""Experiment: one population member with its logger and checkpoint I/O. Same constructor, methods and state dict as
the reference (/root/reference/trainer/experiment.py:16-147). `self.agent` is a view of one slot of a device
population: the Trainer passes `population=` / `slot=` so all experiments share ONE lockstep population; built on its
own (tune_alpha.py --single_experiment) the experiment owns a population of one and `train()` steps it with the
on-device gather + CUDA-graph replay.
"""
from __future__ import annotations

import pickle
from copy import deepcopy
from dataclasses import asdict
from datetime import datetime
from pathlib import Path
from typing import Optional

from evaluator.evaluation import evaluate_agent
from fql.agents.fql import FQLAgent, from_state_dict, to_state_dict
from task.task import Task
from trainer.config import ExperimentConfig, TrainerConfig
from utils.logger import Logger


def resolve_agent_config(trainer_config: TrainerConfig, experiment_config: ExperimentConfig):
    """AgentConfig with the experiment's non-None fields written over it (reference lines 32-41)."""
    tuned = [(f, v) for f, v in vars(experiment_config).items() if v is not None]
    agent_config = deepcopy(trainer_config.agent)
    for f, v in tuned:
        if hasattr(agent_config, f):
            setattr(agent_config, f, v)
    return agent_config, tuned


class Experiment:
    def __init__(self, task: Task, trainer_config: TrainerConfig, experiment_config: ExperimentConfig,
                 state_dict: Optional[dict] = None, population=None, slot: int = 0):
        agent_config, tuned = resolve_agent_config(trainer_config, experiment_config)
        example_batch = task.sample("train", 1)
        self.agent = FQLAgent.create(agent_config.seed, example_batch["observations"], example_batch["actions"],
                                     asdict(agent_config), population=population, slot=slot,
                                     device=getattr(task, "device", None))
        if state_dict is None:
            name = "_".join(f"{f}_{v}" for f, v in tuned)
            self.experiment_name = f"{name}_{datetime.now().strftime('%Y%m%d_%H%M%S')}"
            self.current_step = 0
        else:
            # "agent": None = the member's device state is already in its slot (Trainer._migrate moved it GPU to GPU)
            if state_dict.get("agent") is not None:
                self.agent = from_state_dict(self.agent, state_dict["agent"])
            if "agent_rng" in state_dict:
                self.agent.rng = state_dict["agent_rng"].copy()
            self.experiment_name = state_dict["experiment_name"]
            self.current_step = state_dict["current_step"]
        self.logger = Logger(trainer_config.save_directory, trainer_config.env_name, self.experiment_name, agent_config,
                             use_wandb=trainer_config.use_wandb, state_dict=state_dict["logger"] if state_dict else None)
        self.task = task
        self.steps = trainer_config.steps
        self.log_interval = trainer_config.log_interval
        self.save_directory = Path(trainer_config.save_directory)
        self.env_name = trainer_config.env_name
        self._bound = None

    def state_dict(self) -> dict:
        return {"experiment_name": self.experiment_name, "logger": self.logger.state_dict(),
                "agent": to_state_dict(self.agent), "current_step": self.current_step}

    # ---- training --------------------------------------------------------------------------------------------
    def _bind_train(self):
        pop = self.agent.population
        if hasattr(self.task, "device_dataset"):
            pop.set_dataset(*self.task.device_dataset("train"))
            pop.set_noise()          # on-device Philox gather + noise
            return True
        return False

    def log_step(self, update_info: dict):
        """Train / validation rows at a log point (reference lines 111-116)."""
        self.logger.log(update_info, step=self.current_step, group="train")
        val_batch = self.task.sample("val", self.agent.config["batch_size"])
        _, val_info = self.agent.total_loss(val_batch, grad_params=None)
        self.logger.log(val_info, step=self.current_step, group="val")
        self._bound = None

    def train(self, num_steps: int) -> bool:
        """Trains this member alone for `num_steps` (capped by the step budget). True when the budget is used up."""
        num_steps = min(num_steps, self.steps - self.current_step)
        pop = self.agent.population
        if pop.active != [self.agent.slot]:
            pop.set_active([self.agent.slot])
        done = 0
        while done < num_steps:
            to_log = self.log_interval - self.current_step % self.log_interval
            chunk = min(num_steps - done, to_log)
            if self._bound is None:
                self._bound = self._bind_train()
            if self._bound:
                pop.step(chunk)
                info = None
            else:  # a Task without device datasets: host batches, one at a time (the reference's own loop)
                for _ in range(chunk):
                    self.agent, info = self.agent.update(self.task.sample("train", self.agent.config["batch_size"]))
            self.current_step += chunk
            done += chunk
            if self.logger and self.current_step % self.log_interval == 0:
                self.log_step(info or pop.info_dict(self.agent.slot, with_grad_stats=True))
        return self.current_step == self.steps

    def evaluate(self) -> float:
        eval_info, _ = evaluate_agent(agent=self.agent, env=self.task)
        if self.logger:
            self.logger.log(eval_info, step=self.current_step, group="eval")
        return eval_info["success"]

    def save_agent(self, checkpoint: bool = False):
        filename = f"checkpoint_{self.current_step}.pkl" if checkpoint else "params.pkl"
        path = self.save_directory / self.env_name / self.experiment_name / filename
        path.parent.mkdir(parents=True, exist_ok=True)
        with open(path, "wb") as f:
            pickle.dump(dict(agent=to_state_dict(self.agent)), f)

    def stop(self):
        self.logger.close()
