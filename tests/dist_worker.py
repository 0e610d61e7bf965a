"""Worker of tests/test_dist_cpu.py: one rank of a world_size-2 gloo job running the Trainer over fake device populations."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (os.path.join(ROOT, "flow-q-learning_b200"), ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def main(rank, world, port, out_dir, mode="plain"):
    import numpy as np
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pathlib import Path

    from fakes import FakePopulation, FakeTask
    from hpo.successive_halving import SuccessiveHalving
    from trainer.config import AgentConfig, ExperimentConfig, TrainerConfig
    from trainer.trainer import Trainer
    cfg = TrainerConfig(steps=40, eval_interval=10, log_interval=5, save_directory=Path(out_dir) / f"rank{rank}",
                        env_name="cube-test", agent=AgentConfig(batch_size=8))
    alphas = np.logspace(np.log10(3), 3, 8).tolist()
    migrate = mode.startswith("migrate")
    if mode in ("static", "migrate", "migrate_resume"):
        # FakeTask ranks by |log10(alpha) - 2|: make the best four all start on rank 0 and the rest on rank 1. Placement
        # follows the iteration order of the strategy's candidate SET, so pick alphas whose set order does that
        # (hashes of float / int dataclasses are the same in every process).
        import itertools
        good = (100.0, 90.0, 120.0, 70.0)
        for alphas in itertools.permutations(good + (3.0, 5.0, 7.0, 1000.0)):
            order = list({ExperimentConfig(seed=i, alpha=a) for i, a in enumerate(alphas)})
            if all((c.alpha in good) == (i % 2 == 0) for i, c in enumerate(order)):
                break
        else:
            raise AssertionError("no ordering puts the best four on rank 0")
    pop = [ExperimentConfig(seed=i, alpha=a) for i, a in enumerate(alphas)]
    sh = SuccessiveHalving(set(pop), total_evaluations=4, fraction=0.5, history_length=1)
    if mode.startswith("churn"):
        # 24 members on `world` ranks, scores that reshuffle every evaluation, only two spare slots per rank: many plans,
        # moves in both directions, re-entries of pruned members, moves skipped for want of a free slot
        from fakes import ChurnTask
        cfg = TrainerConfig(steps=160, eval_interval=10, log_interval=5, save_directory=Path(out_dir) / f"rank{rank}",
                            env_name="cube-test", agent=AgentConfig(batch_size=8))
        pop = [ExperimentConfig(seed=i, alpha=a) for i, a in enumerate(np.logspace(np.log10(3), 3, 24).tolist())]
        sh = SuccessiveHalving(set(pop), total_evaluations=16, fraction=0.25, history_length=1)
        tr = Trainer(ChurnTask(), sh, cfg, population_factory=FakePopulation, rebalance=mode == "churn_migrate", spare_slots=2)
    else:
        tr = Trainer(FakeTask(), sh, cfg, population_factory=FakePopulation, rebalance=migrate)
    if mode.endswith("resume"):
        # stop after the first sweep + half of the second, write this rank's shard, rebuild everything from it
        import pickle
        tr.train(max_evaluations=10)
        shard = pickle.loads(pickle.dumps({"trainer": tr.state_dict(), "strategy": tr.strategy.state_dict()}))
        sh = SuccessiveHalving(set(pop), total_evaluations=4, fraction=0.5, history_length=1, state_dict=shard["strategy"])
        tr = Trainer(FakeTask(), sh, cfg, state_dict=shard["trainer"], population_factory=FakePopulation,
                     rebalance=migrate)
    tr.train(max_evaluations=100)
    out = {"rank": rank, "local_members": len(tr.experiments), "candidates": sorted(c.alpha for c in tr.candidates),
           "scores": {str(c.alpha): v for c, v in sh.candidate_scores.items()},
           "steps": {str(c.alpha): e.current_step for c, e in tr.experiments.items()},
           "migrations": tr.migrations, "owner": {str(c.alpha): r for c, r in tr._owner.items()},
           "slots": {str(c.alpha): s for c, s in tr._slot.items()}, "free_count": list(tr._free_count),
           "capacity": list(tr._capacity), "free_slots": sorted(tr._free_slots),
           "imports": [c for c in (tr.population.calls if tr.population else []) if c[0] == "import"],
           "device_steps": {str(c.alpha): int(tr.population.steps_done[tr._slot[c]]) for c in tr.experiments}}
    with open(os.path.join(out_dir, f"rank{rank}.json"), "w") as f:
        json.dump(out, f)
    dist.destroy_process_group()


if __name__ == "__main__":
    main(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], *(sys.argv[5:6]))
