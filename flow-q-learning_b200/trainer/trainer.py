"""Trainer: the reference's population loop (/root/reference/trainer/trainer.py:11-120) -- train every candidate for
`eval_interval` steps, evaluate, `strategy.update`, `strategy.sample`, create / stop experiments, `max_evaluations`
budget, `state_dict` -- with the candidate loop turned into ONE lockstep device population:

  reference                                   here
  for config in candidates: train(...)   ->   population.set_active(slots); population.step(n)   (all at once)
  experiment.evaluate() per candidate    ->   task.rollout_population(population)                (one fused launch)
  strategy.update(experiment, score)     ->   strategy.update(config, score)   (keyed by config, as
                                              /root/reference/evaluator/strategy_evaluator.py:55 does; the reference
                                              trainer passes the Experiment object, which breaks SuccessiveHalving)

Multi-GPU (one process per GPU, torch.distributed initialised by the launcher): candidate i of the initial population
starts on rank i mod world_size -- members never communicate while training -- and the only collective of a sweep is the
all-gather of the evaluation scores, after which every rank runs the same strategy update.

Survivor migration (the single-process reference has no counterpart): when `strategy.sample()` returns a new
candidate set, every rank computes the same plan that evens out the number of candidates per GPU (`_plan_moves`), and
each move ships ONE member -- parameters, Adam moments, target critic, counters, alpha, seed, Philox position: one
device blob, `Population.export_member` / `import_member` -- from its GPU into a spare slot of another with one NCCL
broadcast on the job's communicator. Pruned candidates stay where they are (the reference's re-entry quirk can bring them back), which is why
every rank's population has a few spare slots. The owner map, the free-slot counts and the plan are replicated
deterministic state: no rank ever asks another where a member lives.
"""
from __future__ import annotations

import math
import os
import random
import time
from dataclasses import asdict
from typing import Callable, Dict, List, Optional

import numpy as np

from hpo.strategy import HpoStrategy
from task.task import Task
from trainer.config import TrainerConfig
from trainer.experiment import Experiment, resolve_agent_config


def _default_population_factory(agent_config: dict, alphas, seeds, device):
    from fqlpop.population import Population  # CUDA library; no CPU fallback
    return Population(agent_config, alphas, seeds, device=device)


def _dist():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist
    except ImportError:
        pass
    return None


class Trainer:
    def __init__(self, task: Task, strategy: HpoStrategy, config: TrainerConfig, state_dict: Optional[dict] = None,
                 population_factory: Optional[Callable] = None, rebalance: Optional[bool] = None,
                 spare_slots: Optional[int] = None):
        """`rebalance` (multi-rank jobs only; default on, FQLPOP_REBALANCE=0 turns it off): migrate survivors so that
        every GPU holds about the same number of live candidates. `spare_slots`: extra population slots per rank that
        receive migrants (default: half of the rank's initial share, at least 2)."""
        self.task = task
        self.strategy = strategy
        self.config = config
        self.experiments: Dict = {}
        dist = _dist()
        self.rank = dist.get_rank() if dist else 0
        self.world = dist.get_world_size() if dist else 1
        self._factory = population_factory or _default_population_factory
        if rebalance is None:
            rebalance = os.environ.get("FQLPOP_REBALANCE", "1") != "0"
        self.rebalance = bool(rebalance) and self.world > 1
        self.migrations = 0          # members this job has moved between GPUs (every rank counts all of them)
        self.migration_seconds = 0.0
        self.migration_breakdown: Dict[str, float] = {}

        if state_dict is not None:
            self.candidates = state_dict["candidates"]
            # "order" = the full initial population in slot order (every rank saves it); checkpoints written before it
            # existed are single-rank ones, where the experiment keys are that list
            initial = list(state_dict.get("order") or state_dict["experiments"].keys())
            saved_world = state_dict.get("world", 1)
            if saved_world != self.world:
                raise ValueError(f"checkpoint was written by {saved_world} rank(s), this job has {self.world}: member "
                                 "placement (candidate i on rank i mod world) would not match the saved shards")
        else:
            self.candidates = self.strategy.sample()
            initial = list(self.candidates)
        # a candidate keeps its position in `_order` for life; its GPU and slot change only through `_migrate`
        self._order = {c: i for i, c in enumerate(initial)}
        share = lambda r: len(range(r, len(initial), self.world))      # initial members of rank r
        if state_dict is not None and "owner" in state_dict:
            self._owner = dict(state_dict["owner"])
            self._slot = dict(state_dict["slots"])
            self._capacity = list(state_dict["capacity"])
        else:
            self._owner = {c: i % self.world for i, c in enumerate(initial)}
            self._slot = {c: s for s, c in enumerate(c for c in initial if self.owns(c))}
            if spare_slots is None:
                spare_slots = max(2, math.ceil(share(0) / 2)) if self.rebalance else 0
            self._capacity = [share(r) + (int(spare_slots) if self.rebalance else 0) for r in range(self.world)]
        used = [sum(1 for c in initial if self._owner[c] == r) for r in range(self.world)]
        self._free_count = [self._capacity[r] - used[r] for r in range(self.world)]       # replicated on every rank
        self._free_slots = sorted(set(range(self._capacity[self.rank])) - set(self._slot.values()))
        mine = [c for c in initial if self.owns(c)]
        self.population = None
        if self._capacity[self.rank] > 0:
            example = task.sample("train", 1)
            base = asdict(resolve_agent_config(config, initial[0])[0])
            base["ob_dims"] = tuple(np.shape(example["observations"])[-1:])
            base["action_dim"] = int(np.shape(example["actions"])[-1])
            # spare slots hold a placeholder member (alpha 1, seed 0) until a migrant overwrites it
            alphas, seeds = [1.0] * self._capacity[self.rank], [0] * self._capacity[self.rank]
            for c in mine:
                cfg = resolve_agent_config(config, c)[0]
                alphas[self._slot[c]], seeds[self._slot[c]] = float(cfg.alpha), int(cfg.seed)
            self.population = self._factory(base, alphas, seeds, getattr(task, "device", None))

        if state_dict is not None:
            for c, st in state_dict["experiments"].items():
                self.create_experiment(c, state_dict=st)
            self.untrained_candidates = state_dict["untrained_candidates"]
            self.finished_candidates = state_dict["finished_candidates"]
            self._remote_steps = dict(state_dict.get("remote_steps", {}))
            random.setstate(state_dict["random_rng_state"])
            np.random.set_state(state_dict["np_rng_state"])
        else:
            self.untrained_candidates = []
            self.finished_candidates = []
            for c in self.candidates:
                self.create_experiment(c)

    # ---- placement -----------------------------------------------------------------------------------------------
    def owns(self, candidate) -> bool:
        return self._owner[candidate] == self.rank

    def _ordered(self, candidates) -> List:
        return sorted(candidates, key=lambda c: self._order[c])

    def state_dict(self) -> dict:
        """This rank's shard: its own experiments plus the replicated bookkeeping every rank needs to resume (the
        reference's keys, /root/reference/trainer/trainer.py:50-59, + order / world / remote_steps)."""
        return {
            "experiments": {c: e.state_dict() for c, e in self.experiments.items()},
            "candidates": self.candidates,
            "order": self._ordered(self._order),
            "world": self.world,
            "owner": dict(self._owner),
            "slots": dict(self._slot),
            "capacity": list(self._capacity),
            "remote_steps": dict(getattr(self, "_remote_steps", {})),
            "random_rng_state": random.getstate(),
            "np_rng_state": np.random.get_state(),
            "untrained_candidates": self.untrained_candidates,
            "finished_candidates": self.finished_candidates,
        }

    def create_experiment(self, experiment_config, **kwargs) -> None:
        if experiment_config not in self._order:
            raise KeyError(f"{experiment_config} is not part of the initial population; slots are fixed at start")
        if self.owns(experiment_config):
            self.experiments[experiment_config] = Experiment(
                self.task, self.config, experiment_config, population=self.population,
                slot=self._slot[experiment_config], **kwargs)

    # ---- lockstep pieces ---------------------------------------------------------------------------------------
    def _train_lockstep(self, mine: List, num_steps: int) -> None:
        groups: Dict = {}
        for c in mine:
            e = self.experiments[c]
            n = min(num_steps, e.steps - e.current_step)
            if n > 0:
                groups.setdefault((e.current_step, n), []).append(c)
        pop = self.population
        for (start, n), members in groups.items():
            slots = [self._slot[c] for c in members]
            pop.set_active(slots)
            on_device = hasattr(self.task, "device_dataset")
            step, end = start, start + n
            log_every = self.config.log_interval
            bound = False
            while step < end:
                chunk = min(end - step, log_every - step % log_every)
                if on_device:
                    if not bound:
                        pop.set_dataset(*self.task.device_dataset("train"))
                        pop.set_noise()
                        bound = True
                    pop.step(chunk)
                else:
                    for _ in range(chunk):
                        self._step_host_batches(members, slots)
                step += chunk
                for c in members:
                    self.experiments[c].current_step = step
                if step % log_every == 0:
                    info = pop.info(with_grad_stats=True)
                    keys = self._info_keys()
                    for c, s in zip(members, slots):
                        self.experiments[c].logger.log({k: float(info[s, i]) for i, k in enumerate(keys)}, step=step,
                                                       group="train")
                    self._log_validation(members, slots, step)
                    bound = False

    def _info_keys(self):
        from fqlpop.population import INFO_KEYS
        return INFO_KEYS

    def _bind_host_batches(self, split: str, slots: List[int]) -> None:
        import torch
        pop = self.population
        B = pop.B
        batches = [self.task.sample(split, B) for _ in slots]
        cat = lambda k: np.concatenate([np.asarray(b[k], np.float32).reshape(B, -1) for b in batches])
        idx = torch.zeros(pop.P, B, dtype=torch.int64)
        for j, s in enumerate(slots):
            idx[s] = torch.arange(j * B, (j + 1) * B)
        pop.set_dataset(cat("observations"), cat("actions"), cat("next_observations"), cat("rewards").reshape(-1),
                        cat("masks").reshape(-1))
        pop.set_noise(indices=idx)

    def _step_host_batches(self, members, slots) -> None:
        self._bind_host_batches("train", slots)
        self.population.step(1)

    def _log_validation(self, members, slots, step) -> None:
        self._bind_host_batches("val", slots)
        self.population.total_loss()
        info = self.population.info()
        for c, s in zip(members, slots):
            self.experiments[c].logger.log({k: float(info[s, i]) for i, k in enumerate(self._info_keys()[:10])},
                                           step=step, group="val")

    def _evaluate(self, mine: List) -> Dict:
        if not mine:
            return {}
        if hasattr(self.task, "rollout_population"):
            slots = [self._slot[c] for c in mine]
            self.population.set_active(slots)
            rate = self.task.rollout_population(self.population)["success_rate"]
            rate = rate.cpu().numpy() if hasattr(rate, "cpu") else np.asarray(rate)
            scores = {c: float(rate[s]) for c, s in zip(mine, slots)}
            for c in mine:
                e = self.experiments[c]
                e.logger.log({"success": scores[c]}, step=e.current_step, group="eval")
            return scores
        return {c: self.experiments[c].evaluate() for c in mine}

    def _gather_scores(self, now: List, local: Dict) -> Dict:
        """All-gather of the evaluation scores (the one collective of the whole job)."""
        dist = _dist()
        if dist is None or self.world == 1:
            return dict(local)
        import torch
        per_rank = [[c for c in now if self._owner[c] == r] for r in range(self.world)]
        width = max(1, max(len(x) for x in per_rank))
        device = "cuda" if dist.get_backend() == "nccl" else "cpu"
        send = torch.full((width,), float("nan"), dtype=torch.float32, device=device)
        for j, c in enumerate(per_rank[self.rank]):
            send[j] = local[c]
        recv = [torch.empty_like(send) for _ in range(self.world)]
        dist.all_gather(recv, send)
        return {c: float(recv[r][j]) for r in range(self.world) for j, c in enumerate(per_rank[r])}

    # ---- the loop (control flow of reference lines 69-120) -------------------------------------------------------
    def train(self, max_evaluations: int) -> None:
        while max_evaluations > 0:
            todo = self._ordered(self.untrained_candidates if len(self.untrained_candidates) > 0 else self.candidates)
            now, later = todo[:max_evaluations], todo[max_evaluations:]
            start_time = time.perf_counter()
            mine = [c for c in now if self.owns(c)]
            if mine:
                self._train_lockstep(mine, self.config.eval_interval)
            scores = self._gather_scores(now, self._evaluate(mine))
            for c in now:
                if self.owns(c):
                    e = self.experiments[c]
                    if e.current_step == e.steps and c not in self.finished_candidates:
                        e.save_agent()
                done_steps = min(self.config.steps, self._steps_of(c))
                if done_steps == self.config.steps and c not in self.finished_candidates:
                    self.finished_candidates.append(c)
                self.strategy.update(c, scores[c])
                max_evaluations -= 1
            print(f"Elapsed time: {time.perf_counter() - start_time:.6f} seconds ({len(now)} candidates in lockstep)")

            if later:
                self.untrained_candidates = later
                break
            self.untrained_candidates = []
            if all(c in self.finished_candidates for c in self.candidates):
                break
            new_candidates = self.strategy.sample()
            if self.rebalance:
                self._migrate(self._plan_moves(new_candidates))
            for c in new_candidates:
                if c not in self.experiments and self.owns(c):
                    self.create_experiment(c)
            for c in self.candidates:
                if c not in new_candidates and c in self.experiments:
                    self.experiments[c].stop()
            self.candidates = new_candidates

        for e in self.experiments.values():
            e.stop()

    # ---- survivor migration ---------------------------------------------------------------------------------------
    def _plan_moves(self, candidates) -> List:
        """[(candidate, from_rank, to_rank)]: moves that even out the live candidates per rank, as few as possible
        (a rank gives one away only while it holds at least two more than the emptiest rank with a free slot). Pure
        function of replicated state, so every rank computes the same list."""
        held = {r: [] for r in range(self.world)}
        for c in self._ordered(candidates):
            held[self._owner[c]].append(c)
        free = list(self._free_count)
        moves = []
        while True:
            src = max(range(self.world), key=lambda r: (len(held[r]), -r))
            open_ranks = [r for r in range(self.world) if r != src and free[r] > 0]
            if not open_ranks:
                break
            dst = min(open_ranks, key=lambda r: (len(held[r]), r))
            if len(held[src]) - len(held[dst]) < 2:
                break
            c = held[src].pop()
            held[dst].append(c)
            free[src] += 1
            free[dst] -= 1
            moves.append((c, src, dst))
        return moves

    def _migrate(self, moves: List) -> None:
        if not moves:
            return
        import torch
        dist = _dist()
        t0 = time.perf_counter()
        # what an Experiment is besides its device state: a few hundred bytes per member, one object all-gather
        leaving = {self._order[c]: {"experiment_name": self.experiments[c].experiment_name,
                                    "logger": self.experiments[c].logger.state_dict(),
                                    "current_step": self.experiments[c].current_step,
                                    "agent": None, "agent_rng": self.experiments[c].agent.rng.copy()}
                   for c, src, _ in moves if src == self.rank}
        everyones = [None] * self.world
        dist.all_gather_object(everyones, leaving)
        t1 = time.perf_counter()
        pop = self.population
        # One broadcast per move on the job's own communicator, into a scratch blob every rank keeps: the owner exports
        # into it, the destination imports from it, the others drop it (NVSwitch: ~0.25 ms per 48 MB member on 2 GPUs).
        # Point-to-point send / recv was measured and dropped: NCCL sets up a new channel per pair and direction on first
        # use -- 0.8 s + 0.14 s on 2 GPUs, more than most rungs save -- while the broadcast costs 42 ms once.
        if getattr(self, "_migrate_buf", None) is None:
            self._migrate_buf = torch.empty(pop.member_state_words(), dtype=torch.float32, device=pop.device)
        buf = self._migrate_buf
        t2 = t1
        for c, src, dst in moves:          # plan order, the same on every rank
            if src == self.rank:
                pop.export_member(self._slot[c], out=buf)
            dist.broadcast(buf, src)       # (stream-ordered after the export, and before the import / the next export)
            if src == self.rank:
                e = self.experiments.pop(c)
                e.stop()
                self._remote_steps = getattr(self, "_remote_steps", {})
                self._remote_steps[c] = e.current_step
                self._free_slots.append(self._slot.pop(c))
            elif dst == self.rank:
                slot = self._free_slots.pop(0)
                pop.import_member(slot, buf)
                self._slot[c] = slot
                self._owner[c] = dst
                self.create_experiment(c, state_dict=everyones[src][self._order[c]])
                getattr(self, "_remote_steps", {}).pop(c, None)
            self._owner[c] = dst
            self._free_count[src] += 1
            self._free_count[dst] -= 1
        if buf.is_cuda:
            torch.cuda.current_stream(buf.device).synchronize()
        t3 = time.perf_counter()
        self.migrations += len(moves)
        self.migration_seconds += t3 - t0
        for k, dt in (("metadata_s", t1 - t0), ("transfer_s", t3 - t2)):
            self.migration_breakdown[k] = self.migration_breakdown.get(k, 0.0) + dt

    def _steps_of(self, candidate) -> int:
        """Training steps a candidate has done; every rank can compute it (lockstep: evaluations x eval_interval)."""
        if candidate in self.experiments:
            return self.experiments[candidate].current_step
        self._remote_steps = getattr(self, "_remote_steps", {})
        self._remote_steps[candidate] = min(self.config.steps,
                                            self._remote_steps.get(candidate, 0) + self.config.eval_interval)
        return self._remote_steps[candidate]
