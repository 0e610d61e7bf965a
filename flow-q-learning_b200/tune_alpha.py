"""alpha sweep entry point: same command line, alpha grid, seeds, checkpoint file and single-experiment mode as the
reference's /root/reference/tune_alpha.py:14-107. The population of (alpha, seed) candidates trains in lockstep on the
device (trainer.Trainer -> libfqlpop.so) instead of one agent after the other; under `torchrun` every rank holds its
share of the members and the ranks only exchange evaluation scores.

Additions over the reference's flags (all optional):
  --simulated_evaluations        evaluate in the learned world model (OfflineTaskWithSimulatedEvaluations; needs trained
                                 env models under --save_directory) -- real ogbench/MuJoCo evaluation is unavailable here
  --model / --max_episode_steps  world-model kind and truncation of the simulated task
  --strategy identity|successive_halving, --fraction, --history_length, --total_evaluations
                                 the reference script hard-codes Identity; evaluate_strategy.py replays SuccessiveHalving
"""
import itertools
import pickle
import random

import numpy as np

from argparser import build_config_from_args, get_argparser
from hpo.identity import Identity
from hpo.successive_halving import SuccessiveHalving
from trainer.config import ExperimentConfig
from trainer.experiment import Experiment
from trainer.trainer import Trainer


def main(argv=None):
    parser = get_argparser()
    parser.add_argument("--max_evaluations", type=int, default=200, help="Maximum number of evaluations.")
    parser.add_argument("--number_of_seeds", type=int, default=1,
                        help="Number of random seeds to use per hyperparameter configuration.")
    parser.add_argument("--number_of_alphas", type=int, default=1, help="Number of alpha values to use.")
    parser.add_argument("--simulated_evaluations", action="store_true")
    parser.add_argument("--model", type=str, default="baseline")
    parser.add_argument("--max_episode_steps", type=int, default=1000)
    parser.add_argument("--strategy", choices=("identity", "successive_halving"), default="identity")
    parser.add_argument("--fraction", type=float, default=0.5)
    parser.add_argument("--history_length", type=int, default=1)
    parser.add_argument("--total_evaluations", type=int, default=0)
    args = parser.parse_args(argv)
    config = build_config_from_args(args)

    random.seed(config.seed)
    np.random.seed(config.seed)
    alpha_values = np.logspace(np.log10(3), np.log10(1000), num=args.number_of_alphas).tolist()
    seeds = random.sample(range(10000), args.number_of_seeds)
    combinations = list(itertools.product(alpha_values, seeds))

    if args.simulated_evaluations:
        from task.offline_task_simulated import OfflineTaskWithSimulatedEvaluations
        task = OfflineTaskWithSimulatedEvaluations(config.env_name, model=args.model, data_directory=config.data_directory,
                                                   save_directory=config.save_directory, buffer_size=config.buffer_size,
                                                   num_evaluation_envs=config.eval_episodes,
                                                   max_episode_steps=args.max_episode_steps)
    else:
        from task.offline_task_real import OfflineTaskWithRealEvaluations
        task = OfflineTaskWithRealEvaluations(config.env_name, buffer_size=config.buffer_size,
                                              data_directory=config.data_directory,
                                              num_evaluation_envs=config.eval_episodes)

    if args.single_experiment:
        alpha, seed = combinations[args.job_id]
        experiment = Experiment(task, config, ExperimentConfig(seed=seed, alpha=alpha))
        done = False
        while not done:
            done = experiment.train(config.eval_interval)
            experiment.save_agent(checkpoint=True)
        experiment.stop()
        return experiment

    experiment_configs = [ExperimentConfig(seed=seed, alpha=alpha) for alpha, seed in combinations]
    state_dict = {}
    checkpoint_path = config.save_directory / config.env_name / "checkpoint.pkl"
    if checkpoint_path.exists():
        with open(checkpoint_path, "rb") as f:
            state_dict = pickle.load(f)
    if args.strategy == "identity":
        strategy = Identity(population=experiment_configs, total_evaluations=args.total_evaluations,
                            state_dict=state_dict.get("strategy"))
    else:
        strategy = SuccessiveHalving(population=experiment_configs, total_evaluations=args.total_evaluations,
                                     fraction=args.fraction, history_length=args.history_length,
                                     state_dict=state_dict.get("strategy"))
    trainer = Trainer(task, strategy, config, state_dict=state_dict.get("trainer"))
    trainer.train(max_evaluations=args.max_evaluations)
    if trainer.rank == 0:
        checkpoint_path.parent.mkdir(parents=True, exist_ok=True)
        with open(checkpoint_path, "wb") as f:
            pickle.dump({"trainer": trainer.state_dict(), "strategy": trainer.strategy.state_dict()}, f)
    return trainer


if __name__ == "__main__":
    main()
