"""FQLAgent as the reference's host code sees it -- `create / update / total_loss / sample_actions / config` and a
flax-style state dict (/root/reference/trainer/experiment.py:44-49,61-63,92,108-115,135;
/root/reference/evaluator/evaluation.py:58-63; /root/reference/utils/agent.py:24-35) -- implemented as a VIEW of one
slot of a device-resident `fqlpop.Population`. All arithmetic runs in libfqlpop.so; there is no CPU fallback.

The fast path never calls `update` per agent: `trainer.Trainer` steps every slot of the population in lockstep. A view
created on its own (`FQLAgent.create(...)` without a population, the `--single_experiment` mode of tune_alpha.py) owns
a population of one.

State-dict layout [UPSTREAM-RECALL: flax.serialization.to_state_dict of seohongpark/fql's FQLAgent; the fork's tree
cannot be inspected here]:
  {"rng": uint32[2], "network": {"step": int, "params": {"modules_critic": {"value_net": {...}},
   "modules_target_critic": {...}, "modules_actor_bc_flow": {"mlp": {...}}, "modules_actor_onestep_flow": {"mlp": {...}}},
   "opt_state": {"0": {"count": int, "mu": <params tree>, "nu": <params tree>}, "1": {}}}}
with leaves Dense_i/{kernel[in,out], bias}, LayerNorm_i/{scale, bias}; the critic's leaves carry a leading ensemble axis.
"""
from __future__ import annotations

import copy
from typing import Dict, Optional

import numpy as np

_MODULES = {"critic": ("modules_critic", "value_net"), "actor_bc_flow": ("modules_actor_bc_flow", "mlp"),
            "actor_onestep_flow": ("modules_actor_onestep_flow", "mlp")}
_TARGET = ("modules_target_critic", "value_net")


def _seed_to_u64(seed) -> int:
    """Accepts an int or a jax-style uint32[2] key."""
    if seed is None:
        return int(np.random.randint(0, 2 ** 31))
    a = np.asarray(seed).astype(np.uint64).ravel()
    if a.size == 1:
        return int(a[0])
    return int((a[0] << np.uint64(32)) | a[1])


class FQLAgent:
    def __init__(self, population, slot: int, config: dict, owns_population: bool = False):
        self.population = population
        self.slot = int(slot)
        self.config = config
        self._owns = owns_population
        self.rng = np.array([0, int(config.get("seed", 0)) & 0xFFFFFFFF], dtype=np.uint32)

    # ---- construction ------------------------------------------------------------------------------------
    @classmethod
    def create(cls, seed, ex_observations, ex_actions, config, population=None, slot: int = 0, device="cuda:0"):
        config = dict(copy.deepcopy(config))
        config["seed"] = int(seed)
        config["ob_dims"] = tuple(np.shape(ex_observations)[-1:])
        config["action_dim"] = int(np.shape(ex_actions)[-1])
        owns = population is None
        if owns:
            from fqlpop.population import Population  # needs a CUDA device; fails loudly otherwise
            population = Population(config, [float(config.get("alpha", 10.0))], [int(seed)], device=device)
        return cls(population, slot, config, owns_population=owns)

    # ---- the three calls on the hot path ---------------------------------------------------------------------
    def _bind_batch(self, batch):
        pop = self.population
        B = len(batch["observations"])
        if B != pop.B:
            raise ValueError(f"batch of {B} rows, population was built for batch_size {pop.B}")
        import torch
        idx = torch.zeros(pop.P, B, dtype=torch.int64)
        idx[self.slot] = torch.arange(B)
        pop.set_dataset(batch["observations"], batch["actions"], batch["next_observations"],
                        np.asarray(batch["rewards"]).reshape(-1), np.asarray(batch["masks"]).reshape(-1))
        pop.set_noise(indices=idx)

    def update(self, batch):
        """One gradient step of THIS member on an explicit host batch -> (agent, info)."""
        pop = self.population
        if pop.active != [self.slot]:
            pop.set_active([self.slot])
        self._bind_batch(batch)
        pop.step(1)
        return self, pop.info_dict(self.slot, with_grad_stats=True)

    def total_loss(self, batch, grad_params=None, rng=None):
        pop = self.population
        if pop.active != [self.slot]:
            pop.set_active([self.slot])
        self._bind_batch(batch)
        pop.total_loss()
        info = pop.info_dict(self.slot)
        return info["critic/critic_loss"] + info["actor/actor_loss"], info

    def sample_actions(self, observations, seed=None, temperature: float = 1.0):
        """clip(pi(obs, z), -1, 1), z ~ N(0, I) from `seed`; `temperature` is accepted and unused, as upstream."""
        obs = np.asarray(observations, dtype=np.float32)
        single = obs.ndim == 1
        out = self.population.sample_actions(self.slot, obs, None, _seed_to_u64(seed)).cpu().numpy()
        return out[0] if single else out

    # ---- (de)serialisation -------------------------------------------------------------------------------------
    def state_dict(self) -> dict:
        pop, m = self.population, self.slot
        params, mu, nu = {}, {}, {}
        for net, (mod, sub) in _MODULES.items():
            params[mod] = {sub: pop.get_tree(m, net, "params")}
            mu[mod] = {sub: pop.get_tree(m, net, "mu")}
            nu[mod] = {sub: pop.get_tree(m, net, "nu")}
        tgt = pop.get_tree(m, "critic", "target")
        params[_TARGET[0]] = {_TARGET[1]: tgt}
        zeros = lambda t: {k: zeros(v) if isinstance(v, dict) else np.zeros_like(v) for k, v in t.items()}
        mu[_TARGET[0]] = {_TARGET[1]: zeros(tgt)}
        nu[_TARGET[0]] = {_TARGET[1]: zeros(tgt)}
        adam_steps, counter = pop.get_counters()
        return {"rng": self.rng.copy(),
                "network": {"step": int(adam_steps[m]) + 1, "params": params,
                            "opt_state": {"0": {"count": int(adam_steps[m]), "mu": mu, "nu": nu}, "1": {}}},
                "fqlpop": {"philox_step": int(counter[m])}}

    def load_state_dict(self, state: dict) -> "FQLAgent":
        pop, m = self.population, self.slot
        net_state = state["network"]
        opt = net_state.get("opt_state", {}).get("0", {})
        for net, (mod, sub) in _MODULES.items():
            pop.set_tree(m, net, net_state["params"][mod][sub], "params")
            if "mu" in opt:
                pop.set_tree(m, net, opt["mu"][mod][sub], "mu")
                pop.set_tree(m, net, opt["nu"][mod][sub], "nu")
        pop.set_tree(m, "critic", net_state["params"][_TARGET[0]][_TARGET[1]], "target")
        pop.refresh_shadow()
        steps, counter = pop.get_counters()
        steps[m] = int(opt.get("count", 0))
        # continue this member's batch-index / noise stream where the saved run stopped (the reference restores the
        # agent rng and the numpy RNG state on resume, /root/reference/trainer/trainer.py:55-56)
        counter[m] = int(state.get("fqlpop", {}).get("philox_step", counter[m]))
        pop.set_counters(adam_steps=steps, step_counter=counter)
        if "rng" in state:
            self.rng = np.asarray(state["rng"], dtype=np.uint32).copy()
        return self

    def close(self):
        if self._owns and self.population is not None:
            self.population.close()
            self.population = None


def to_state_dict(agent: FQLAgent) -> dict:
    """flax.serialization.to_state_dict stand-in used by Experiment / load_agent."""
    return agent.state_dict()


def from_state_dict(agent: FQLAgent, state: dict) -> FQLAgent:
    return agent.load_state_dict(state)
