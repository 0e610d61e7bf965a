"""Configuration of a world-model training run. Field names and defaults are the interface
(/root/reference/envmodel/config.py:5-20: `TrainerConfig`), because argparser.build_env_model_config_from_args fills
it by keyword from the command line and the trainers read it by attribute.
The class is generated from the table below; construct it with keywords."""
from dataclasses import field, make_dataclass
from pathlib import Path
from typing import Any, Dict

# (name, type, default)                                      what reads it
_FIELDS = [
    ("env_name", str, "cube-single-play-singletask-task2-v0"),  # dataset / output directory
    ("model", str, "baseline"),                                  # baseline | multistep | termination_predictor
    ("model_config", Dict[str, Any], dict),                      # --model.* flags (hidden_dims, latent_dim), saved as yaml
    ("steps", int, 2000),                                        # optimiser steps = cosine decay length
    ("batch_size", int, 256),
    ("sequence_length", int, 256),                               # window of the multistep model
    ("init_learning_rate", float, 1e-3),
    ("true_termination_weight", float, 30.0),                    # weight of the positive class in the termination BCE
    ("termination_weight", float, 1.0),                          # weight of the termination term in the state loss
    ("reconstruction_weight", float, 1.0),
    ("val_batches", int, 20),                                    # validation batches per validation pass
    ("seed", int, 0),
    ("data_directory", Path, Path("data/")),
    ("save_directory", Path, Path("exp/")),
]

TrainerConfig = make_dataclass(
    "TrainerConfig",
    [(n, t, field(default_factory=d) if d is dict else field(default=d)) for n, t, d in _FIELDS],
    module=__name__)
