"""Command-line surface of tune_alpha.py / reproduce.py: the same flags, dotted `--agent.*` namespace, defaults and
store_true behaviour as the reference (/root/reference/argparser.py:9-169). Note `--agent.layer_norm` is a store_true
flag, so from the CLI it is False unless passed even though AgentConfig defaults to True -- the reference's launch
scripts always pass it. The world-model training parser of the reference (lines 172-292) is out of scope."""
import argparse
from ast import literal_eval
from dataclasses import fields
from pathlib import Path

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
