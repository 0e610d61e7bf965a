"""alpha sweep entry point. Command line, alpha grid (log-spaced 3..1000), seed draw, (alpha, seed) product, job-array
mode and checkpoint file follow the reference's /root/reference/tune_alpha.py:14-107; the candidates train in lockstep on
the device (Trainer -> libfqlpop.so) and, under `torchrun`, are spread over the ranks. Optional flags of this build
(simulated evaluations, successive halving): see _entry.sweep_parser."""
import itertools
import random

import numpy as np

from _entry import make_task, run_one, run_population, seeded_config, sweep_parser
from trainer.config import ExperimentConfig

_FLAGS = [
    ("--max_evaluations", int, 200, "Maximum number of evaluations."),
    ("--number_of_seeds", int, 1, "Number of random seeds to use per hyperparameter configuration."),
    ("--number_of_alphas", int, 1, "Number of alpha values to use."),
]


def main(argv=None):
    args = sweep_parser(_FLAGS).parse_args(argv)
    config = seeded_config(args)
    alphas = np.logspace(np.log10(3), np.log10(1000), num=args.number_of_alphas).tolist()
    seeds = random.sample(range(10000), args.number_of_seeds)
    grid = list(itertools.product(alphas, seeds))
    task = make_task(args, config)
    if args.single_experiment:
        alpha, seed = grid[args.job_id]
        return run_one(task, config, ExperimentConfig(seed=seed, alpha=alpha),
                       after_round=lambda e: e.save_agent(checkpoint=True))
    return run_population(args, task, config, [ExperimentConfig(seed=s, alpha=a) for a, s in grid])


if __name__ == "__main__":
    main()
