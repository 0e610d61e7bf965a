"""Worker of tests/test_dist_gpu.py: one rank (torchrun) -- or the single-process reference run -- of the real Trainer +
SuccessiveHalving + OfflineTaskWithSimulatedEvaluations.rollout_population over device populations, scores all-gathered
over NCCL. Usage: [torchrun --nproc-per-node N] dist_gpu_worker.py <out_dir> [n_alphas [resume]]"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (os.path.join(ROOT, "flow-q-learning_b200"), ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)
os.environ.setdefault("FQLPOP_SYNTHETIC", "1")
os.environ.setdefault("FQLPOP_SYNTHETIC_ROWS", "20000")
ENV = "cube-single-play-singletask-task2-v0"


def main(out_dir, n_alphas=8, resume=0):
    from pathlib import Path

    import numpy as np
    import torch
    import yaml
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    from _entry import init_distributed
    created = init_distributed()
    from oracle import fql_oracle as orc   # checker-side helper: deterministic world-model weights
    from utils import flax_msgpack
    save = Path(out_dir) / f"w{world}r{rank}"
    root = save / ENV / "env_models"
    root.mkdir(parents=True, exist_ok=True)
    env = orc.init_envmodel(11, 28, 5, hidden=(128, 256, 128))
    env["termination"]["Dense_3"]["bias"] -= 0.5
    to_np = lambda t: orc.tree_map(lambda x: x.numpy(), t)
    for name, key in (("baseline", "state"), ("termination_predictor", "termination")):
        (root / f"{name}.pt").write_bytes(flax_msgpack.to_bytes({"params": to_np(env[key])}))
        with open(root / f"{name}_config.yaml", "w") as f:
            yaml.dump({"hidden_dims": (128, 256, 128)}, f)

    import random
    from hpo.successive_halving import SuccessiveHalving
    from task.offline_task_simulated import OfflineTaskWithSimulatedEvaluations
    from trainer.config import AgentConfig, ExperimentConfig, TrainerConfig
    from trainer.trainer import Trainer
    random.seed(0); np.random.seed(0)
    task = OfflineTaskWithSimulatedEvaluations(ENV, "baseline", data_directory=None, save_directory=save, buffer_size=1000,
                                               num_evaluation_envs=64, max_episode_steps=30)
    cfg = TrainerConfig(steps=12, eval_interval=4, log_interval=4, save_directory=save / "runs", env_name=ENV,
                        agent=AgentConfig(batch_size=64))
    pop = [ExperimentConfig(seed=s, alpha=a) for s, a in enumerate(np.logspace(np.log10(3), 3, n_alphas).tolist())]
    sh = SuccessiveHalving(set(pop), total_evaluations=3, fraction=0.5, history_length=1)
    tr = Trainer(task, sh, cfg)
    assert tr.world == world and tr.rank == rank
    if world == 2:
        # survivor migration over NCCL, exercised whatever the scores turn out to be: ship one of rank 0's members to
        # rank 1 before training starts (3 / 5); the first rebalancing after a rung sends members the other way
        first = next(c for c in tr._ordered(tr._order) if tr._owner[c] == 0)
        tr._migrate([(first, 0, 1)])
        assert tr._owner[first] == 1 and (first in tr.experiments) == (rank == 1)
    t0 = time.perf_counter()
    if resume:
        # stop after the first sweep + a part of the second (i.e. after the first rebalancing), write this rank's shard,
        # throw everything away and rebuild population, experiments and strategy from the shard
        import pickle
        tr.train(max_evaluations=n_alphas + 2)
        shard = pickle.loads(pickle.dumps({"trainer": tr.state_dict(), "strategy": sh.state_dict()}))
        moved = tr.migrations
        if tr.population is not None:
            tr.population.close()
        sh = SuccessiveHalving(set(pop), total_evaluations=3, fraction=0.5, history_length=1, state_dict=shard["strategy"])
        tr = Trainer(task, sh, cfg, state_dict=shard["trainer"])
        tr.migrations = moved
    tr.train(max_evaluations=1000)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    out = {"rank": rank, "world": world, "device": torch.cuda.current_device(),
           "local_members": len(tr.experiments), "candidates": sorted(c.alpha for c in tr.candidates),
           "scores": {repr(c.alpha): [float(x) for x in v] for c, v in sh.candidate_scores.items()},
           "steps": {repr(c.alpha): e.current_step for c, e in tr.experiments.items()}, "seconds": dt,
           "backend": (torch.distributed.get_backend() if world > 1 else None),
           "migrations": tr.migrations, "migration_seconds": tr.migration_seconds}
    with open(os.path.join(out_dir, f"w{world}r{rank}{'_resumed' if resume else ''}.json"), "w") as f:
        json.dump(out, f)
    if tr.population is not None:
        tr.population.close()
    if created:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main(sys.argv[1], *(int(x) for x in sys.argv[2:4]))
