"""load_model(save_directory, env_name, model, sample_batch) -> apply_fn, as the reference
(/root/reference/utils/envmodel.py:9-56): reads `<save>/<env>/env_models/<model>_config.yaml` and `<model>.pt`
(flax msgpack bytes), rebuilds the network, returns a callable.

  baseline / multistep : apply_fn(observations, actions) -> (next_observations, observations)
                         (BaselineStatePredictor.__call__, /root/reference/envmodel/baseline.py:25-37; `multistep`
                          uses the scan cell's weights, params["params"]["ScanCell_0"]["cell"])
  termination_predictor: apply_fn(observations, train=False, rng=None) -> logits
                         (TerminationPredictor.__call__, /root/reference/envmodel/termination_predictor.py:14-21;
                          Dropout is the identity at evaluation)

The returned object also carries `.params` (numpy tree) so OfflineTaskWithSimulatedEvaluations can hand BOTH networks to
one fused device world model. Arithmetic runs in libfqlpop.so (fqlpop_env_step); a network loaded on its own is paired
with an all-zero partner, which the residual connection turns into the identity.
"""
from pathlib import Path
from typing import Dict

import numpy as np
import yaml

from utils.flax_msgpack import from_bytes, to_bytes

_STATE_MODELS = ("baseline", "multistep")


def _zero_mlp(in_dim: int, out_dim: int, with_ln: bool) -> Dict:
    tree = {"Dense_0": {"kernel": np.zeros((in_dim, 64), np.float32), "bias": np.zeros(64, np.float32)},
            "Dense_1": {"kernel": np.zeros((64, out_dim), np.float32), "bias": np.zeros(out_dim, np.float32)}}
    if with_ln:
        tree["LayerNorm_0"] = {"scale": np.ones(in_dim, np.float32), "bias": np.zeros(in_dim, np.float32)}
    return tree


class LoadedModel:
    def __init__(self, kind: str, params: Dict, ob_dim: int, action_dim: int, hidden_dims, device=None):
        self.kind, self.params, self.ob_dim, self.action_dim = kind, params, ob_dim, action_dim
        self.hidden_dims = tuple(hidden_dims)
        if device is None:
            from fqlpop._lib import default_device
            device = default_device()
        self.device = device
        self._dev = None

    def _device_model(self):
        if self._dev is None:
            from fqlpop.envmodel import EnvModel  # needs CUDA; no CPU fallback
            O, A = self.ob_dim, self.action_dim
            if self.kind in _STATE_MODELS:
                self._dev = EnvModel(self.params, _zero_mlp(O, 1, False), O, A, device=self.device)
            else:
                self._dev = EnvModel(_zero_mlp(O + A, O, True), self.params, O, A, device=self.device)
        return self._dev

    def __call__(self, observations, actions=None, train: bool = False, rng=None):
        import torch
        if train:
            raise NotImplementedError("world-model training (Dropout active) is outside the rollout hot path")
        obs = torch.as_tensor(np.array(observations, dtype=np.float32))
        if self.kind in _STATE_MODELS:
            nxt, _ = self._device_model().step(obs, torch.as_tensor(np.array(actions, dtype=np.float32)))
            return nxt.cpu().numpy(), np.asarray(observations)
        zeros = torch.zeros(obs.shape[0], self.action_dim)
        _, logits = self._device_model().step(obs, zeros)
        return logits.cpu().numpy()


def load_model(save_directory: Path, env_name: str, model: str, sample_batch: Dict[str, np.ndarray], device=None):
    root = Path(save_directory) / env_name / "env_models"
    config_path = root / f"{model}_config.yaml"
    if not config_path.exists():
        raise FileNotFoundError(
            f"Configuration file not found at {config_path}. Please ensure the configuration has been set.")
    with open(config_path, "r") as f:
        model_config = yaml.unsafe_load(f)
    if model not in _STATE_MODELS + ("termination_predictor",):
        raise ValueError(f"Unknown model type: {model}")
    params_path = root / f"{model}.pt"
    if not params_path.exists():
        raise FileNotFoundError(f"Model file not found at {params_path}. Please ensure the model has been trained.")
    tree = from_bytes(params_path.read_bytes())
    params = tree["params"]["ScanCell_0"]["cell"] if model == "multistep" else tree["params"]
    return LoadedModel(model, params, int(np.shape(sample_batch["observations"])[-1]),
                       int(np.shape(sample_batch["actions"])[-1]), model_config["hidden_dims"], device=device)


def save_model(params: Dict, save_directory: Path, env_name: str, model: str, model_config: Dict) -> Path:
    """Writes what train_env_model.py leaves behind (/root/reference/train_env_model.py:118-127):
    `<save>/<env>/env_models/<model>_config.yaml` (yaml of model_config) and `<model>.pt` = flax msgpack bytes of
    `trainer.state.params` ({"params": {...}}). Returns the .pt path."""
    root = Path(save_directory) / env_name / "env_models"
    root.mkdir(parents=True, exist_ok=True)
    with open(root / f"{model}_config.yaml", "w") as f:
        yaml.dump(dict(model_config), f)
    tree = params if "params" in params else {"params": params}
    path = root / f"{model}.pt"
    path.write_bytes(to_bytes(tree))
    return path
