"""Command-line surface of tune_alpha.py / reproduce.py: the same flags, dotted `--agent.*` namespace, defaults and
store_true behaviour as the reference (/root/reference/argparser.py:9-169). Note `--agent.layer_norm` is a store_true
flag, so from the CLI it is False unless passed even though AgentConfig defaults to True -- the reference's launch
scripts always pass it. `get_env_model_argparser` / `build_env_model_config_from_args` mirror the world-model training
parser (lines 172-292): `--model.*` flags land in TrainerConfig.model_config."""
import argparse
from ast import literal_eval
from dataclasses import fields
from pathlib import Path

from envmodel.config import TrainerConfig as EnvModelTrainerConfig
from trainer.config import AgentConfig, TrainerConfig

_TRAINER_FLAGS = [
    ("seed", int, 0, "Random seed for the trainer."),
    ("steps", int, 1000000, "Number of training steps."),
    ("log_interval", int, 5000, "Logging interval."),
    ("eval_interval", int, 100000, "Evaluation interval."),
    ("save_directory", str, "exp/", "Directory to save checkpoints and logs."),
    ("data_directory", str, "data/", "Dataset directory."),
    ("env_name", str, "cube-single-play-singletask-task2-v0", "Environment (dataset) name."),
    ("eval_episodes", int, 50, "Number of evaluation episodes."),
    ("buffer_size", int, 2000000, "Replay buffer size."),
]
_AGENT_FLAGS = [
    ("seed", int, 0, "Random seed for the agent."),
    ("agent_name", str, "fql", "Agent name."),
    ("ob_dims", int, None, "Observation dimensions (set automatically)."),
    ("action_dim", int, None, "Action dimension (set automatically)."),
    ("lr", float, 3e-4, "Learning rate."),
    ("batch_size", int, 256, "Batch size."),
    ("actor_hidden_dims", literal_eval, "(512, 512, 512, 512)", "Actor network hidden dimensions (tuple)."),
    ("value_hidden_dims", literal_eval, "(512, 512, 512, 512)", "Value network hidden dimensions (tuple)."),
    ("discount", float, 0.99, "Discount factor."),
    ("tau", float, 0.005, "Target network update rate."),
    ("q_agg", str, "mean", "Aggregation method for target Q values."),
    ("alpha", float, 10.0, "BC coefficient (needs tuning for each environment)."),
    ("flow_steps", int, 10, "Number of flow steps."),
    ("encoder", str, None, "Visual encoder name (e.g., impala_small)."),
]
_AGENT_SWITCHES = [
    ("layer_norm", "Use layer normalization."),
    ("actor_layer_norm", "Use layer normalization in actor."),
    ("normalize_q_loss", "Whether to normalize the Q loss."),
]


def get_argparser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="Trainer and Agent Configuration")
    for name, typ, default, text in _TRAINER_FLAGS:
        parser.add_argument(f"--{name}", type=typ, default=default, help=text)
    parser.add_argument("--use_wandb", action="store_true", help="Use Weights & Biases logging.")
    for name, typ, default, text in _AGENT_FLAGS:
        parser.add_argument(f"--agent.{name}", type=typ, default=default, help=text)
    for name, text in _AGENT_SWITCHES:
        parser.add_argument(f"--agent.{name}", action="store_true", help=text)
    parser.add_argument("--single_experiment", action="store_true", help="Whether to run a single experiment.")
    parser.add_argument("--job_id", type=int, default=0, help="Job ID in a Job Array.")
    return parser


def build_config_from_args(args: argparse.Namespace) -> TrainerConfig:
    agent_names = {f.name for f in fields(AgentConfig)}
    trainer_names = {f.name for f in fields(TrainerConfig)} - {"agent"}
    agent_kwargs, trainer_kwargs = {}, {}
    for key, value in vars(args).items():
        if key.startswith("agent."):
            if key[6:] in agent_names:
                agent_kwargs[key[6:]] = value
        elif key in trainer_names:
            trainer_kwargs[key] = value
    for key in ("save_directory", "data_directory"):
        trainer_kwargs[key] = Path(trainer_kwargs[key])
    if isinstance(agent_kwargs.get("actor_hidden_dims"), str):  # argparse does not run `type` on string defaults here
        agent_kwargs["actor_hidden_dims"] = literal_eval(agent_kwargs["actor_hidden_dims"])
    if isinstance(agent_kwargs.get("value_hidden_dims"), str):
        agent_kwargs["value_hidden_dims"] = literal_eval(agent_kwargs["value_hidden_dims"])
    return TrainerConfig(**trainer_kwargs, agent=AgentConfig(**agent_kwargs))


_ENV_MODEL_FLAGS = [
    ("model", str, "baseline", "The environment model to be used."),
    ("model.hidden_dims", literal_eval, (128, 256, 128), "The dimension of hidden layers."),
    ("model.latent_dim", int, 4, "The dimension of latent representation."),
    ("true_termination_weight", float, 30.0, "Weight for the true class in the termination loss in the env model."),
    ("termination_weight", float, 1.0, "Weight for the termination loss in the env model."),
    ("reconstruction_weight", float, 1.0, "Weight for the reconstruction loss in the env model."),
    ("sequence_length", int, 256, "The length of sequences for the multistep model."),
    ("steps", int, 20000, "The number of training steps."),
    ("env_name", str, "cube-single-play-singletask-task2-v0", "The environment task."),
    ("init_learning_rate", float, 1e-3, "Initial learning rate."),
    ("seed", int, 0, "Random seed for training."),
    ("batch_size", int, 256, "Batch size."),
    ("val_batches", int, 50, "Number of validation batches."),
    ("data_directory", str, "data/", "Path to a directory where training data is stored."),
    ("save_directory", str, "exp/", "Path to a directory where model parameters will be saved."),
]


def get_env_model_argparser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="The configurations of the environment model.")
    for name, typ, default, text in _ENV_MODEL_FLAGS:
        parser.add_argument(f"--{name}", type=typ, default=default, help=text)
    return parser


def build_env_model_config_from_args(args: argparse.Namespace) -> EnvModelTrainerConfig:
    model_config, trainer_kwargs = {}, {}
    for key, value in vars(args).items():
        if key.startswith("model."):
            model_config[key[len("model."):]] = value
        else:
            trainer_kwargs[key] = value
    for key in ("save_directory", "data_directory"):
        trainer_kwargs[key] = Path(trainer_kwargs[key])
    names = {f.name for f in fields(EnvModelTrainerConfig)}
    return EnvModelTrainerConfig(**{k: v for k, v in trainer_kwargs.items() if k in names}, model_config=model_config)
