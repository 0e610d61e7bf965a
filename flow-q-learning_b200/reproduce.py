"""Seed sweep at the configured alpha: same command line, seeds, checkpoint file and single-experiment mode as the
reference's /root/reference/reproduce.py:11-82 (one ExperimentConfig(seed=...) per seed, Identity strategy). The members
train in lockstep on the device; `--simulated_evaluations` (+ `--model`, `--max_episode_steps`) evaluates in the learned
world model, since real ogbench/MuJoCo evaluation is unavailable here."""
import pickle
import random

import numpy as np

from argparser import build_config_from_args, get_argparser
from hpo.identity import Identity
from trainer.config import ExperimentConfig
from trainer.experiment import Experiment
from trainer.trainer import Trainer


def main(argv=None):
    parser = get_argparser()
    parser.add_argument("--max_evaluations", type=int, default=200, help="Maximum number of evaluations.")
    parser.add_argument("--number_of_seeds", type=int, default=1, help="Number of random seeds to use.")
    parser.add_argument("--simulated_evaluations", action="store_true")
    parser.add_argument("--model", type=str, default="baseline")
    parser.add_argument("--max_episode_steps", type=int, default=1000)
    args = parser.parse_args(argv)
    config = build_config_from_args(args)
    random.seed(config.seed)
    np.random.seed(config.seed)
    seeds = random.sample(range(10000), args.number_of_seeds)
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
        experiment = Experiment(task, config, ExperimentConfig(seed=seeds[args.job_id]))
        done = False
        while not done:
            done = experiment.train(config.eval_interval)
            experiment.evaluate()
        experiment.save_agent()
        experiment.stop()
        return experiment
    experiment_configs = [ExperimentConfig(seed=seed) for seed in seeds]
    state_dict = {}
    checkpoint_path = config.save_directory / config.env_name / "checkpoint.pkl"
    if checkpoint_path.exists():
        with open(checkpoint_path, "rb") as f:
            state_dict = pickle.load(f)
    strategy = Identity(population=experiment_configs, total_evaluations=0, state_dict=state_dict.get("strategy"))
    trainer = Trainer(task, strategy, config, state_dict=state_dict.get("trainer"))
    trainer.train(max_evaluations=args.max_evaluations)
    if trainer.rank == 0:
        checkpoint_path.parent.mkdir(parents=True, exist_ok=True)
        with open(checkpoint_path, "wb") as f:
            pickle.dump({"trainer": trainer.state_dict(), "strategy": trainer.strategy.state_dict()}, f)
    return trainer


if __name__ == "__main__":
    main()
