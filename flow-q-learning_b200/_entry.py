"""Shared plumbing of the sweep entry points (tune_alpha.py, reproduce.py): extra command-line flags, task construction,
checkpoint file handling and the two run modes of the reference scripts (one experiment of a job array, or the whole
population under a Trainer)."""
import pickle
import random
from typing import Callable, List, Sequence

import numpy as np

from argparser import build_config_from_args, get_argparser
from hpo.identity import Identity
from hpo.successive_halving import SuccessiveHalving
from trainer.config import ExperimentConfig
from trainer.experiment import Experiment
from trainer.trainer import Trainer


def sweep_parser(extra: Sequence[tuple]):
    """The reference's parser plus `extra` = (flag, type, default, help) and this build's optional flags."""
    parser = get_argparser()
    for flag, typ, default, text in extra:
        parser.add_argument(flag, type=typ, default=default, help=text)
    group = parser.add_argument_group("this build")
    group.add_argument("--simulated_evaluations", action="store_true",
                       help="Evaluate in the learned world model (real ogbench/MuJoCo evaluation is unavailable here).")
    group.add_argument("--model", type=str, default="baseline", help="World-model kind of the simulated task.")
    group.add_argument("--max_episode_steps", type=int, default=1000, help="Truncation of simulated episodes.")
    group.add_argument("--strategy", choices=("identity", "successive_halving"), default="identity")
    group.add_argument("--fraction", type=float, default=0.5)
    group.add_argument("--history_length", type=int, default=1)
    group.add_argument("--total_evaluations", type=int, default=0)
    return parser


def seeded_config(args):
    config = build_config_from_args(args)
    random.seed(config.seed)
    np.random.seed(config.seed)
    return config


def make_task(args, config):
    common = dict(buffer_size=config.buffer_size, data_directory=config.data_directory,
                  num_evaluation_envs=config.eval_episodes)
    if args.simulated_evaluations:
        from task.offline_task_simulated import OfflineTaskWithSimulatedEvaluations
        return OfflineTaskWithSimulatedEvaluations(config.env_name, model=args.model, save_directory=config.save_directory,
                                                   max_episode_steps=args.max_episode_steps, **common)
    from task.offline_task_real import OfflineTaskWithRealEvaluations
    return OfflineTaskWithRealEvaluations(config.env_name, **common)


def run_one(task, config, experiment_config: ExperimentConfig, after_round: Callable[[Experiment], None], at_end=None):
    """single_experiment mode: train in rounds of eval_interval until the step budget is spent."""
    experiment = Experiment(task, config, experiment_config)
    while not experiment.train(config.eval_interval):
        after_round(experiment)
    after_round(experiment)
    if at_end is not None:
        at_end(experiment)
    experiment.stop()
    return experiment


def run_population(args, task, config, population: List[ExperimentConfig]):
    """Population mode: resume from <save>/<env>/checkpoint.pkl if present, train, write the checkpoint back (rank 0)."""
    path = config.save_directory / config.env_name / "checkpoint.pkl"
    saved = pickle.loads(path.read_bytes()) if path.exists() else {}
    kw = dict(population=population, total_evaluations=args.total_evaluations, state_dict=saved.get("strategy"))
    strategy = (SuccessiveHalving(fraction=args.fraction, history_length=args.history_length, **kw)
                if args.strategy == "successive_halving" else Identity(**kw))
    trainer = Trainer(task, strategy, config, state_dict=saved.get("trainer"))
    trainer.train(max_evaluations=args.max_evaluations)
    if trainer.rank == 0:
        path.parent.mkdir(parents=True, exist_ok=True)
        path.write_bytes(pickle.dumps({"trainer": trainer.state_dict(), "strategy": trainer.strategy.state_dict()}))
    return trainer
