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
    group.add_argument("--generated_initial_observations", action="store_true",
                       help="Simulated episodes start from samples of the trained initial-observation VAE "
                            "(train_env_model.py --model initial_observation) instead of environment resets.")
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
                                                   max_episode_steps=args.max_episode_steps,
                                                   generated_initial_observations=args.generated_initial_observations,
                                                   **common)
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


def init_distributed():
    """One process per GPU under torchrun: torchrun only sets RANK / WORLD_SIZE / LOCAL_RANK, so the process group the
    Trainer shards over is created here (NCCL on GPUs, gloo without). Returns True when this call created the group."""
    import os
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return False
    import torch
    import torch.distributed as dist
    if dist.is_initialized():
        return False
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if torch.cuda.is_available():
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group("gloo")
    return True


def checkpoint_path(config, rank: int):
    """Rank 0 keeps the reference's file name; the other ranks write their shard (their members' weights and optimiser
    state) beside it."""
    name = "checkpoint.pkl" if rank == 0 else f"checkpoint.rank{rank}.pkl"
    return config.save_directory / config.env_name / name


def run_population(args, task, config, population: List[ExperimentConfig]):
    """Population mode: resume from <save>/<env>/checkpoint[.rankR].pkl if present, train, write this rank's shard back."""
    created = init_distributed()
    try:
        return _run_population(args, task, config, population)
    finally:
        if created:
            import torch.distributed as dist
            dist.destroy_process_group()


def _run_population(args, task, config, population: List[ExperimentConfig]):
    import os
    rank = int(os.environ.get("RANK", "0")) if int(os.environ.get("WORLD_SIZE", "1")) > 1 else 0
    path = checkpoint_path(config, rank)
    saved = pickle.loads(path.read_bytes()) if path.exists() else {}
    kw = dict(population=population, total_evaluations=args.total_evaluations, state_dict=saved.get("strategy"))
    strategy = (SuccessiveHalving(fraction=args.fraction, history_length=args.history_length, **kw)
                if args.strategy == "successive_halving" else Identity(**kw))
    trainer = Trainer(task, strategy, config, state_dict=saved.get("trainer"))
    trainer.train(max_evaluations=args.max_evaluations)
    path = checkpoint_path(config, trainer.rank)
    path.parent.mkdir(parents=True, exist_ok=True)
    path.write_bytes(pickle.dumps({"trainer": trainer.state_dict(), "strategy": trainer.strategy.state_dict()}))
    return trainer
