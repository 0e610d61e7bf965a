"""Initial-observation generator: the decoder half of the reference's VAE
(/root/reference/envmodel/initial_observation.py:21-52) loaded from the files the reference's analysis script reads
(/root/reference/distribution_visualizer.py:61-94): `<save>/<env>/env_models/initial_observation_config.yaml`
(`hidden_dims`, `latent_dim`) and `initial_observation.pt` (flax msgpack; the decoder is ["params"]["decoder"], built with
hidden_dims[::-1]). Sampling runs in libfqlpop.so (fqlpop_decoder_sample)."""
from dataclasses import dataclass
from pathlib import Path
from typing import Tuple

import yaml

from utils.flax_msgpack import from_bytes


@dataclass(frozen=True)
class Decoder:
    output_dimension: int
    hidden_dims: Tuple[int, ...]


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
