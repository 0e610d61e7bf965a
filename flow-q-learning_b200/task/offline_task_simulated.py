"""OfflineTaskWithSimulatedEvaluations -- same constructor, `sample / reset / step / close` as the reference
(/root/reference/task/offline_task_simulated.py:15-111), with the world-model arithmetic on the GPU:

  step(actions)            one transition of every episode through libfqlpop.so (fqlpop_env_step): state predictor,
                           termination predictor, reward = where(term, 0, -1), truncation at max_episode_steps, and the
                           `invalid` flag from the step after an episode first ended (reference lines 85-107);
  rollout_population(pop)  the fast path: ALL episodes of ALL active population members for the whole evaluation in one
                           fused persistent kernel (fqlpop_rollout) -- policy, world model and the bookkeeping of
                           evaluate_actor_fn (/root/reference/evaluator/evaluation.py:67-114) without leaving the GPU;
  device_dataset(split)    the offline dataset resident in HBM for the on-device batch gather.

Initial observations: the reference resets REAL MuJoCo environments (lines 60-83); those are unavailable on this box, so
the reset-only environments of fql.envs.env_utils return dataset episode-start observations instead.
"""
from __future__ import annotations

from pathlib import Path
from typing import Optional

import numpy as np

from fql.envs.env_utils import make_env_and_datasets
from fql.utils.datasets import Dataset, ReplayBuffer
from task.task import Task
from utils.envmodel import load_model


class OfflineTaskWithSimulatedEvaluations(Task):
    def __init__(self, env_name: str, model: str = "baseline", data_directory: Path = None,
                 save_directory: Path = Path("exp/"), buffer_size: int = 2000000, num_evaluation_envs: int = 50,
                 max_episode_steps: int = 1000, device=None, generated_initial_observations: bool = False):
        self.max_episode_steps = max_episode_steps
        self.num_evaluation_envs = num_evaluation_envs
        if device is None:
            from fqlpop._lib import default_device
            device = default_device()
        self.device = device
        _, self.eval_envs, train, val = make_env_and_datasets(env_name, data_directory,
                                                              num_evaluation_envs=num_evaluation_envs)
        if num_evaluation_envs == 1:
            self.eval_envs = [self.eval_envs]
        self.train_dataset = Dataset.create(**train)
        self.train_dataset = ReplayBuffer.create_from_initial_dataset(
            dict(self.train_dataset), size=max(buffer_size, self.train_dataset.size + 1))
        self.val_dataset = val if isinstance(val, Dataset) else Dataset.create(**val)
        example = self.train_dataset.sample(1)
        self.state_predictor = load_model(save_directory, env_name, model, example, device=device)
        self.termination_predictor = load_model(save_directory, env_name, "termination_predictor", example, device=device)
        # optional: initial observations from the trained VAE decoder on the device instead of environment resets
        self.initial_observation_decoder = None
        if generated_initial_observations:
            from envmodel.initial_observation import load_initial_observation_decoder
            self.initial_observation_decoder = load_initial_observation_decoder(
                save_directory, env_name, self.state_predictor.ob_dim, device=device)
        self.current_observations = None
        self.episode_steps = 0
        self.invalidate = []
        self.model = model
        self.env_name = env_name
        self._fused = None
        self._device_sets = {}

    # ---- data ---------------------------------------------------------------------------------------------------
    def sample(self, dataset, batch_size: int):
        return (self.train_dataset if dataset == "train" else self.val_dataset).sample(batch_size)

    def device_dataset(self, split: str = "train"):
        if split not in self._device_sets:
            ds = self.train_dataset if split == "train" else self.val_dataset
            self._device_sets[split] = ds.device_arrays(self.device)
        return self._device_sets[split]

    # ---- world model --------------------------------------------------------------------------------------------
    def envmodel(self):
        """Both predictors as ONE device world model (what the fused kernels take)."""
        if self._fused is None:
            from fqlpop.envmodel import EnvModel
            sp = self.state_predictor
            self._fused = EnvModel(sp.params, self.termination_predictor.params, sp.ob_dim, sp.action_dim,
                                   device=self.device)
        return self._fused

    def reset(self, seed: Optional[int] = None):
        if len(self.eval_envs) == 1:
            obs, info = self.eval_envs[0].reset(seed=seed)
            observations, infos = [obs], [info]
        elif seed is None:
            observations, infos = zip(*[e.reset() for e in self.eval_envs])
        else:
            observations, infos = zip(*[e.reset(seed=seed + i) for i, e in enumerate(self.eval_envs)])
        self.current_observations = np.array(observations, dtype=np.float32)
        self.episode_steps = 0
        self.invalidate = [False] * len(self.eval_envs)
        return self.current_observations, list(infos)

    def step(self, actions):
        import torch
        nxt, logits = self.envmodel().step(torch.as_tensor(np.asarray(self.current_observations, np.float32)),
                                           torch.as_tensor(np.asarray(actions, np.float32)))
        next_observations = nxt.cpu().numpy()
        terminations = logits.cpu().numpy() > 0
        reward = np.where(terminations, 0, -1)
        self.current_observations = next_observations            # every episode advances, finished ones too
        self.episode_steps += 1
        truncations = np.full_like(terminations, self.episode_steps >= self.max_episode_steps, dtype=bool)
        infos = [{"success": bool(s)} for s in terminations]
        for i in range(len(self.eval_envs)):
            if self.invalidate[i]:
                infos[i]["invalid"] = True
            if terminations[i] or truncations[i]:
                self.invalidate[i] = True
        return next_observations, reward, terminations, truncations, infos

    def rollout_population(self, population, seed: Optional[int] = None, trajectory: bool = False):
        """Simulated evaluation of every active member at once -> dict of device tensors indexed by member id
        (`success_rate[P]` is evaluate_actor_fn's stats["success"])."""
        eval_seed = int(np.random.randint(0, 2 ** 31)) if seed is None else int(seed)
        if self.initial_observation_decoder is not None:
            observations = self.initial_observation_decoder.sample(self.num_evaluation_envs, seed=eval_seed)
        else:
            observations, _ = self.reset(seed=seed)
        return population.rollout(self.envmodel(), observations, self.max_episode_steps, eval_seed=eval_seed,
                                  trajectory=trajectory)

    def close(self):
        for e in self.eval_envs:
            e.close()
