"""Seed sweep at the configured alpha (the reference's /root/reference/reproduce.py:11-82): one candidate per seed,
job-array mode evaluates after every round and saves the agent once at the end. Same plumbing as tune_alpha.py."""
import random

from _entry import make_task, run_one, run_population, seeded_config, sweep_parser
from trainer.config import ExperimentConfig

_FLAGS = [
    ("--max_evaluations", int, 200, "Maximum number of evaluations."),
    ("--number_of_seeds", int, 1, "Number of random seeds to use."),
]


def main(argv=None):
    args = sweep_parser(_FLAGS).parse_args(argv)
    config = seeded_config(args)
    seeds = random.sample(range(10000), args.number_of_seeds)
    task = make_task(args, config)
    if args.single_experiment:
        return run_one(task, config, ExperimentConfig(seed=seeds[args.job_id]), after_round=lambda e: e.evaluate(),
                       at_end=lambda e: e.save_agent())
    return run_population(args, task, config, [ExperimentConfig(seed=s) for s in seeds])


if __name__ == "__main__":
    main()
