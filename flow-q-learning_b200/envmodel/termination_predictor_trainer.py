"""TerminationPredictorTrainer with the reference's constructor and train() loop
(/root/reference/envmodel/termination_predictor_trainer.py:21-171); train_step / eval_step run on the device.

nn.Dropout(0.1) on the input: the reference passes the SAME rng (PRNGKey(config.seed)) to every train_step (:69), so the
dropout mask is one fixed [batch, ob_dim] pattern for the whole run; it is drawn once here (numpy stream, not threefry)
and handed to the kernels as a multiplier keep / 0.9.
"""
from functools import partial
from typing import Dict, Optional

import numpy as np

from envmodel.config import TrainerConfig
from envmodel.loss import focal_loss
from envmodel.state_predictor_trainer import _State
from utils.data_loader import DataLoader


class TerminationPredictorTrainer:
    def __init__(self, model, train_loader: DataLoader, val_loader: DataLoader, loss_fn, config: TrainerConfig,
                 logger=None, device=None, device_sampling: bool = False, params: Optional[Dict] = None):
        from fqlpop.worldmodel import TERMINATION, WorldModelTrainer  # needs libfqlpop.so; no CPU fallback
        self.model, self.train_loader, self.val_loader = model, train_loader, val_loader
        self.loss_fn, self.config, self.logger = loss_fn, config, logger
        self.device_sampling = device_sampling
        fn = loss_fn.func if isinstance(loss_fn, partial) else loss_fn
        if getattr(fn, "__name__", "") != focal_loss.__name__:
            raise ValueError("TerminationPredictorTrainer: the device path implements envmodel.loss.focal_loss "
                             "(the loss train_env_model.py:81 selects)")
        kw = dict(loss_fn.keywords) if isinstance(loss_fn, partial) else {}
        sample_batch = self.train_loader.sample(self.config.batch_size)
        ob_dim = int(np.shape(sample_batch["observations"])[-1])
        rate = float(getattr(model, "dropout_rate", 0.1))
        keep = np.random.default_rng(config.seed).random((config.batch_size, ob_dim)) >= rate
        self.dropout_mask = (keep / (1.0 - rate)).astype(np.float32)
        self.device_trainer = WorldModelTrainer(
            TERMINATION, ob_dim, int(np.shape(sample_batch["actions"])[-1]), tuple(model.hidden_dims), config.batch_size,
            config.steps, config.init_learning_rate, params=params, seed=config.seed,
            focal_alpha=float(kw.get("alpha", 0.25)), focal_gamma=float(kw.get("gamma", 2.0)), device=device)
        self.state = _State(self)

    def schedule(self, step: int) -> float:
        return self.config.init_learning_rate * 0.5 * (1 + np.cos(np.pi * min(step, self.config.steps) / self.config.steps))

    def train_step(self, batch: Dict[str, np.ndarray]) -> Dict[str, float]:
        return self.device_trainer.step(batch, train=True, dropout_mask=self.dropout_mask)

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
            self.device_trainer.set_dataset(self.train_loader.dataset)
        while step < self.config.steps:
            if step % 100 == 0:
                self._validate(step)
            if self.device_sampling:
                n = min(100 - step % 100, self.config.steps - step)
                self.device_trainer.train(n, dropout_mask=self.dropout_mask)
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
