"""Initial-observation generator: the reference's VAE (/root/reference/envmodel/initial_observation.py:9-78).

* The model classes carry the reference's names and fields (Encoder, Decoder, InitialObservationEnvModel); they are
  descriptions -- the arithmetic runs in libfqlpop.so.
* `vae_loss` has the reference's name; like the losses of envmodel/loss.py it is what a trainer is configured with
  (functools.partial(vae_loss, reconstruction_weight=...)), the device kernel vae_loss_kernel computes it; the numpy body
  serves host-side analysis.
* `InitialObservationTrainer`: the reference ships the model and the loss but no trainer (train_env_model.py does not
  know the model, although argparser.py:190-196,212-216 already carries --model.latent_dim and
  --reconstruction_weight). This one has the constructor / train() loop / logging cadence of StatePredictorTrainer
  (/root/reference/envmodel/state_predictor_trainer.py:22-170) with train_step / eval_step on the device
  (fqlpop_vae_step / fqlpop_vae_train), and leaves the files the reference's analysis script reads
  (/root/reference/distribution_visualizer.py:61-94): `<save>/<env>/env_models/initial_observation_config.yaml`
  (`hidden_dims`, `latent_dim`) and `initial_observation.pt` (flax msgpack of {"params": {"encoder", "decoder"}}; the
  decoder is built with hidden_dims[::-1]).
* `load_initial_observation_decoder` reads those files back; sampling runs in libfqlpop.so (fqlpop_decoder_sample)."""
from dataclasses import dataclass
from functools import partial
from pathlib import Path
from typing import Dict, Optional, Tuple

import numpy as np
import yaml

from utils.flax_msgpack import from_bytes


@dataclass(frozen=True)
class Encoder:
    latent_dimension: int
    hidden_dims: Tuple[int, ...]


@dataclass(frozen=True)
class Decoder:
    output_dimension: int
    hidden_dims: Tuple[int, ...]


@dataclass(frozen=True)
class InitialObservationEnvModel:
    observation_dimension: int
    latent_dimension: int
    hidden_dims: Tuple[int, ...] = (128,)

    @property
    def encoder(self) -> Encoder:
        return Encoder(self.latent_dimension, tuple(self.hidden_dims))

    @property
    def decoder(self) -> Decoder:
        return Decoder(self.observation_dimension, tuple(self.hidden_dims)[::-1])


def vae_loss(observations, reconstructed_observations, mu, logvar, reconstruction_weight: float):
    """The arithmetic of the reference's vae_loss (:55-78) on arrays (the reference applies the model inside)."""
    observations = np.asarray(observations, np.float32)
    rl = np.mean((observations - np.asarray(reconstructed_observations, np.float32)) ** 2)
    mu, logvar = np.asarray(mu, np.float32), np.asarray(logvar, np.float32)
    kl = -0.5 * np.mean(1 + logvar - mu ** 2 - np.exp(logvar))
    loss = (reconstruction_weight * rl + kl) / (reconstruction_weight + 1)
    return loss, {"reconstruction_loss": rl, "kl": kl, "loss": loss}


def load_initial_observation_decoder(save_directory: Path, env_name: str, observation_dimension: int, device=None):
    root = Path(save_directory) / env_name / "env_models"
    config_path = root / "initial_observation_config.yaml"
    if not config_path.exists():
        raise FileNotFoundError(
            f"Configuration file not found at {config_path}. Please ensure the configuration has been set.")
    with open(config_path, "r") as f:
        model_config = yaml.unsafe_load(f)
    model_path = root / "initial_observation.pt"
    if not model_path.exists():
        raise FileNotFoundError(f"Model file not found at {model_path}. Please ensure the model has been trained.")
    params = from_bytes(model_path.read_bytes())["params"]["decoder"]
    from fqlpop.worldmodel import Decoder as DeviceDecoder  # needs libfqlpop.so; no CPU fallback
    return DeviceDecoder(params, model_config["latent_dim"], tuple(model_config["hidden_dims"])[::-1],
                         observation_dimension, device=device)


class _VaeState:
    """Stand-in for flax's TrainState: `.params` is what flax.serialization.to_bytes receives."""

    def __init__(self, trainer):
        self._t = trainer

    @property
    def params(self) -> Dict:
        return {"params": self._t.device_trainer.params()}

    @property
    def step(self) -> int:
        return int(self._t.device_trainer.logs()[11])


class InitialObservationTrainer:
    def __init__(self, model: InitialObservationEnvModel, train_loader, val_loader, loss_fn, config, logger=None,
                 device=None, device_sampling: bool = False, params: Optional[Dict] = None):
        from fqlpop.worldmodel import VaeTrainer  # needs libfqlpop.so; no CPU fallback
        fn = loss_fn.func if isinstance(loss_fn, partial) else loss_fn
        if getattr(fn, "__name__", "") != vae_loss.__name__:
            raise ValueError("InitialObservationTrainer: the device path implements envmodel.initial_observation.vae_loss only")
        kw = dict(loss_fn.keywords) if isinstance(loss_fn, partial) else {}
        self.reconstruction_weight = float(kw.get("reconstruction_weight", config.reconstruction_weight))
        self.model, self.train_loader, self.val_loader = model, train_loader, val_loader
        self.loss_fn, self.config, self.logger = loss_fn, config, logger
        self.device_sampling = device_sampling
        self.device_trainer = VaeTrainer(model.observation_dimension, model.latent_dimension, tuple(model.hidden_dims),
                                         config.batch_size, config.steps, config.init_learning_rate,
                                         self.reconstruction_weight, params=params, seed=config.seed, device=device)
        self.state = _VaeState(self)

    def schedule(self, step: int) -> float:
        return self.config.init_learning_rate * 0.5 * (1 + np.cos(np.pi * min(step, self.config.steps) / self.config.steps))

    def train_step(self, batch: Dict[str, np.ndarray], eps=None) -> Dict[str, float]:
        return self.device_trainer.step(batch["observations"], eps, train=True)

    def eval_step(self, batch: Dict[str, np.ndarray], eps=None) -> Dict[str, float]:
        return self.device_trainer.step(batch["observations"], eps, train=False)

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
            self.device_trainer.set_dataset(self.train_loader.dataset["observations"])
        while step < self.config.steps:
            if step % 100 == 0:
                self._validate(step)
            if self.device_sampling:
                n = min(100 - step % 100, self.config.steps - step)
                self.device_trainer.train(n)
                logs, step = self.device_trainer.named_logs(), step + n
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
