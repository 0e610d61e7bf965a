"""make_env_and_datasets with the fork's signature as the reference calls it
(/root/reference/task/offline_task_real.py:26-31: `(env_name, data_directory, num_evaluation_envs=n)` ->
`(env, eval_envs, train_dataset, val_dataset)`).

ogbench / MuJoCo are not installed here ("unavailable on this box"), so:
  * datasets come from `<data_directory>/<env_name>.npz` and `<env_name>-val.npz` when those files exist (OGBench's
    on-disk layout: observations, actions, terminals [, next_observations, rewards, masks]); a missing file raises
    FileNotFoundError. ONLY with the explicit opt-in FQLPOP_SYNTHETIC=1 (tests, benchmarks) a synthetic dataset with the
    OGBench shape of that environment family is generated instead (N rows, obs ~ N(0,1),
    next = obs + 0.05 N(0,1), actions ~ U(-1,1), rewards in {-1,0} with P(0)=0.005, 1000-step episodes);
  * evaluation "environments" only know how to reset: they return dataset episode-start observations (the same rows
    InitialObservationLoader keeps, /root/reference/utils/data_loader.py:14-22). Their dynamics are the learned world
    model on the GPU (task/offline_task_simulated.py); calling step() on them raises.
"""
from __future__ import annotations

import os
from pathlib import Path
from typing import Optional

import numpy as np

ENV_SHAPES = {"cube-single": (28, 5), "cube-double": (37, 5), "antsoccer": (42, 8)}
EPISODE_LENGTH = 1000
SYNTHETIC_ROWS = int(os.environ.get("FQLPOP_SYNTHETIC_ROWS", "1000000"))


def env_shape(env_name: str):
    for prefix, shape in ENV_SHAPES.items():
        if env_name.startswith(prefix):
            return shape
    raise ValueError(f"unknown environment family for {env_name!r}; known: {sorted(ENV_SHAPES)}")


def synthetic_dataset(env_name: str, rows: int, seed: int = 0) -> dict:
    O, A = env_shape(env_name)
    rng = np.random.default_rng(seed)
    obs = rng.standard_normal((rows, O), dtype=np.float32)
    nxt = obs + 0.05 * rng.standard_normal((rows, O), dtype=np.float32)
    act = np.clip(rng.uniform(-1, 1, (rows, A)).astype(np.float32), -1 + 1e-5, 1 - 1e-5)
    rew = -(rng.random(rows) >= 0.005).astype(np.float32)
    terminals = np.zeros(rows, np.float32)
    terminals[EPISODE_LENGTH - 1:: EPISODE_LENGTH] = 1.0
    return {"observations": obs, "actions": act, "next_observations": nxt, "rewards": rew,
            "masks": (rew != 0).astype(np.float32), "terminals": terminals}


def _load_npz(path: Path) -> dict:
    with np.load(path) as f:
        d = {k: f[k] for k in f.files}
    obs = d["observations"].astype(np.float32)
    terminals = d["terminals"].astype(np.float32)
    if "next_observations" not in d:  # OGBench compact layout: next row of the same episode
        keep = np.nonzero(terminals == 0)[0]
        keep = keep[keep + 1 < len(obs)]  # a trailing row without terminals == 1 has no successor
        d = {"observations": obs[keep], "actions": d["actions"][keep].astype(np.float32),
             "next_observations": obs[keep + 1], "terminals": terminals[keep + 1],
             **({"rewards": d["rewards"][keep].astype(np.float32)} if "rewards" in d else {})}
    if "rewards" not in d:
        d["rewards"] = -np.ones(len(d["observations"]), np.float32)
    if "masks" not in d:
        d["masks"] = (d["rewards"] != 0).astype(np.float32)
    return {k: np.asarray(v, dtype=np.float32) for k, v in d.items()}


class ResetOnlyEnv:
    """Stand-in for one evaluation environment: reset(seed) -> an episode-start observation of the dataset."""

    def __init__(self, initial_observations: np.ndarray, index: int):
        self._init = initial_observations
        self._index = index

    def reset(self, seed: Optional[int] = None, **_):
        k = self._index if seed is None else int(seed)
        return self._init[k % len(self._init)].copy(), {}

    def step(self, action):
        raise RuntimeError("real environment dynamics are unavailable on this box; use the learned world model")

    def close(self):
        pass


def synthetic_allowed() -> bool:
    return os.environ.get("FQLPOP_SYNTHETIC", "0") not in ("", "0")


def make_env_and_datasets(env_name: str, data_directory=None, num_evaluation_envs: int = 1, **_):
    """Returns (env, eval_envs, train_dataset, val_dataset). A missing dataset file raises; generated data is used only
    under FQLPOP_SYNTHETIC=1 and announced on stdout."""
    train = val = None
    p = Path(data_directory) / f"{env_name}.npz" if data_directory is not None else None
    if p is not None and p.exists():
        train = _load_npz(p)
        pv = Path(data_directory) / f"{env_name}-val.npz"
        if pv.exists():
            val = _load_npz(pv)
        elif not synthetic_allowed():
            raise FileNotFoundError(f"validation dataset {pv} not found (set FQLPOP_SYNTHETIC=1 to train on generated data)")
    elif not synthetic_allowed():
        raise FileNotFoundError(
            f"dataset {p if p is not None else '<data_directory>/' + env_name + '.npz'} not found: pass --data_directory "
            "with the OGBench .npz files. Agents and world models are NOT trained on generated data unless "
            "FQLPOP_SYNTHETIC=1 is set explicitly (tests and benchmarks do).")
    if train is None:
        print(f"[fqlpop] data: synthetic ({SYNTHETIC_ROWS} generated rows of {env_name} shape; FQLPOP_SYNTHETIC=1)")
        train = synthetic_dataset(env_name, SYNTHETIC_ROWS, seed=0)
    if val is None:
        val = synthetic_dataset(env_name, max(EPISODE_LENGTH, len(train["observations"]) // 10), seed=1)
    starts = train["observations"][::EPISODE_LENGTH]
    envs = [ResetOnlyEnv(starts, i) for i in range(num_evaluation_envs)]
    eval_envs = envs[0] if num_evaluation_envs == 1 else envs
    return envs[0], eval_envs, train, val
