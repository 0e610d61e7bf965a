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
    """{"a": {"b": 1}} -> {"a.b": 1} (nested info dicts of an environment step)."""
    flat, stack = {}, [(parent_key, d)]
    while stack:
        prefix, node = stack.pop()
        for name, value in node.items():
            path = name if not prefix else f"{prefix}{sep}{name}"
            if hasattr(value, "items"):
                stack.append((path, value))
            else:
                flat[path] = value
    return flat


def add_to(dict_of_lists, single_dict):
    for name in single_dict:
        dict_of_lists[name].append(single_dict[name])


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
    """Host loop over any Task: all environments step together until every one of them has been done once; an
    environment contributes its info dict exactly once, at its first done step (later steps carry info["invalid"])."""
    if seed is not None:
        random.seed(seed)
        np.random.seed(seed)
    obs, _ = env.reset(seed=seed)
    n_envs = len(obs)
    finished = np.zeros(n_envs, dtype=bool)
    collected, transitions = defaultdict(list), []
    while not finished.all():
        act = np.clip(np.array(actor_fn(observations=obs, temperature=eval_temperature)), -1, 1)
        nxt, _, terminated, truncated, infos = env.step(act)
        stale = np.fromiter((bool(i.get("invalid", False)) for i in infos), dtype=bool, count=n_envs)
        finished = np.asarray(terminated, dtype=bool) | np.asarray(truncated, dtype=bool) | stale
        for idx in np.flatnonzero(finished & ~stale):
            add_to(collected, flatten(infos[idx]))
        transitions.append((obs, act, stale))
        obs = nxt
    return {name: np.mean(values) for name, values in collected.items()}, transitions
