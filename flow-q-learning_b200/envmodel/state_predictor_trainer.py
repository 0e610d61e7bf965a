"""StatePredictorTrainer with the reference's constructor and train() loop
(/root/reference/envmodel/state_predictor_trainer.py:22-170); train_step / eval_step run on the device
(fqlpop.worldmodel -> fqlpop_wm_step / fqlpop_wm_train in libfqlpop.so).

Loop semantics kept: validation on `val_batches` batches every 100 steps and once more after the last step, logged as
val/<key> means; train/<key> + train/learning_rate every step; one `train_loader.sample(batch_size)` per step, so numpy's
global RNG is consumed exactly as in the reference. `device_sampling=True` instead keeps the training set resident in HBM
and draws batches on the device (Philox), 100 optimiser steps per host call.
"""
from functools import partial
from typing import Dict, Optional

import numpy as np

from envmodel.config import TrainerConfig
from envmodel.loss import state_prediction_loss
from utils.data_loader import DataLoader
from utils.envmodel import load_model


class _State:
    """Stand-in for flax's TrainState: `.params` is what flax.serialization.to_bytes receives (train_env_model.py:126)."""

    def __init__(self, trainer):
        self._t = trainer

    @property
    def params(self) -> Dict:
        tree = self._t.device_trainer.params()
        if getattr(self._t, "multistep", False):  # nn.scan(Cell) with the baseline predictor as `cell` (multistep.py:17-21,41-53)
            tree = {"ScanCell_0": {"cell": tree}}
        return {"params": tree}

    @property
    def step(self) -> int:
        return int(self._t.device_trainer.logs()[11])


def _loss_weights(loss_fn, config: TrainerConfig):
    fn = loss_fn.func if isinstance(loss_fn, partial) else loss_fn
    if getattr(fn, "__name__", "") != state_prediction_loss.__name__:
        raise ValueError("StatePredictorTrainer: the device path implements envmodel.loss.state_prediction_loss only")
    kw = dict(loss_fn.keywords) if isinstance(loss_fn, partial) else {}
    return (float(kw.get("true_termination_weight", config.true_termination_weight)),
            float(kw.get("termination_weight", config.termination_weight)),
            float(kw.get("reconstruction_weight", config.reconstruction_weight)))


class StatePredictorTrainer:
    def __init__(self, model, train_loader: DataLoader, val_loader: DataLoader, loss_fn, config: TrainerConfig,
                 logger=None, device=None, device_sampling: bool = False, params: Optional[Dict] = None):
        from fqlpop.worldmodel import MULTISTEP, STATE, WorldModelTrainer  # needs libfqlpop.so; no CPU fallback
        from envmodel.multistep import MultistepStatePredictor
        self.multistep = isinstance(model, MultistepStatePredictor)
        self.model, self.train_loader, self.val_loader = model, train_loader, val_loader
        self.loss_fn, self.config, self.logger = loss_fn, config, logger
        self.device_sampling = device_sampling
        sample_batch = self.train_loader.sample(self.config.batch_size)  # the reference draws one batch for init (:42)
        ttw, tw, rw = _loss_weights(loss_fn, config)
        term_params, term_hidden = None, ()
        if tw > 0:  # frozen, already trained (state_predictor_trainer.py:47-56)
            tp = load_model(config.save_directory, config.env_name, "termination_predictor",
                            {k: np.reshape(v, (-1,) + np.shape(v)[2:]) if self.multistep else v for k, v in sample_batch.items()},
                            device=device)
            term_params, term_hidden = tp.params, tp.hidden_dims
        self.device_trainer = WorldModelTrainer(
            MULTISTEP if self.multistep else STATE, int(np.shape(sample_batch["observations"])[-1]),
            int(np.shape(sample_batch["actions"])[-1]),
            tuple(model.hidden_dims), config.batch_size, config.steps, config.init_learning_rate, params=params,
            seed=config.seed, termination_params=term_params, termination_hidden_dims=term_hidden,
            true_termination_weight=ttw, termination_weight=tw, reconstruction_weight=rw,
            sequence_length=int(np.shape(sample_batch["actions"])[1]) if self.multistep else 1, device=device)
        self.state = _State(self)

    def schedule(self, step: int) -> float:
        return self.config.init_learning_rate * 0.5 * (1 + np.cos(np.pi * min(step, self.config.steps) / self.config.steps))

    def train_step(self, batch: Dict[str, np.ndarray]) -> Dict[str, float]:
        return self.device_trainer.step(batch, train=True)

    def eval_step(self, batch: Dict[str, np.ndarray]) -> Dict[str, float]:
        return self.device_trainer.step(batch, train=False)

    def _validate(self, step: int):
        val_logs = [self.eval_step(self.val_loader.sample(self.config.batch_size)) for _ in range(self.config.val_batches)]
        means = {k: float(np.mean(np.array([log[k] for log in val_logs], dtype=np.float32))) for k in val_logs[0]}
        if self.logger:
            for k, v in means.items():
                self.logger.log({f"val/{k}": v}, step=step)
        return means

    def train(self) -> None:
        step = 0
        if self.device_sampling:
            data = self.train_loader.dataset
            if self.multistep:  # MultistepLoader keeps [episodes, 1000, ...]; the device sampler wants the flat rows back
                data = {k: np.reshape(v, (-1,) + tuple(np.shape(v)[2:])) for k, v in data.items()}
            self.device_trainer.set_dataset(data)
        while step < self.config.steps:
            if step % 100 == 0:
                self._validate(step)
            if self.device_sampling:
                n = min(100 - step % 100, self.config.steps - step)
                self.device_trainer.train(n)
                logs, step = self.device_trainer.named_logs(train=True), step + n
                if self.logger:
                    self.logger.log({"train/learning_rate": self.schedule(step - 1),
                                     **{f"train/{k}": v for k, v in logs.items()}}, step=step - 1)
                continue
            logs = self.train_step(self.train_loader.sample(self.config.batch_size))
            if self.logger:
                self.logger.log({"train/learning_rate": self.schedule(step), **{f"train/{k}": v for k, v in logs.items()}},
                                step=step)
            step += 1
        self.last_validation = self._validate(max(step - 1, 0))
