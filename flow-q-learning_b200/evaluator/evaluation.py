"""evaluate_agent / evaluate_actor_fn with the reference's signatures and statistics
(/root/reference/evaluator/evaluation.py:41-114): success = mean over environments of `terminated` at the
environment's FIRST done step; finished environments keep being stepped but are masked by `invalid`.

When `env` offers `rollout_population` (the simulated task) and `agent` is a population view, the whole evaluation is one
fused kernel launch; otherwise the generic host loop below drives any Task step by step.
"""
from __future__ import annotations

import random
from collections import defaultdict
from typing import Callable, Optional

import numpy as np


def supply_rng(f, rng=0):
    """Gives every call of f a fresh `seed=` (a counter-derived 64-bit value; the reference splits a jax key)."""
    state = {"key": int(np.asarray(rng).astype(np.uint64).ravel()[-1])}

    def wrapped(*args, **kwargs):
        state["key"] = (state["key"] * 6364136223846793005 + 1442695040888963407) % (1 << 64)
        return f(*args, seed=state["key"], **kwargs)

    return wrapped


def flatten(d, parent_key="", sep="."):
    out = {}
    for k, v in d.items():
        key = f"{parent_key}{sep}{k}" if parent_key else k
        if hasattr(v, "items"):
            out.update(flatten(v, key, sep=sep))
        else:
            out[key] = v
    return out


def add_to(dict_of_lists, single_dict):
    for k, v in single_dict.items():
        dict_of_lists[k].append(v)


def evaluate_agent(agent, env, seed: Optional[int] = None, eval_temperature: float = 0):
    population = getattr(agent, "population", None)
    if population is not None and hasattr(env, "rollout_population"):
        if population.active != [agent.slot]:
            population.set_active([agent.slot])
        out = env.rollout_population(population, seed=seed)
        return {"success": float(out["success_rate"][agent.slot].item())}, []
    actor_fn = supply_rng(agent.sample_actions, rng=seed if seed is not None else np.random.randint(0, 2 ** 32))
    return evaluate_actor_fn(actor_fn, env, seed, eval_temperature)


def evaluate_actor_fn(actor_fn: Callable, env, seed: Optional[int] = None, eval_temperature: float = 0):
    if seed is not None:
        random.seed(seed)
        np.random.seed(seed)
    stats = defaultdict(list)
    observations, infos = env.reset(seed=seed)
    done = np.zeros(len(observations), dtype=bool)
    transitions = []
    while not np.all(done):
        actions = np.clip(np.array(actor_fn(observations=observations, temperature=eval_temperature)), -1, 1)
        next_observations, _, terminated, truncated, infos = env.step(actions)
        mask = np.array([info.get("invalid", False) for info in infos], dtype=bool)
        done = np.logical_or(np.logical_or(terminated, truncated), mask)
        for i, info in enumerate(infos):
            if done[i] and not mask[i]:      # the environment's first done step
                add_to(stats, flatten(info))
        transitions.append((observations, actions, mask))
        observations = next_observations
    return {k: np.mean(v) for k, v in stats.items()}, transitions
