"""CPU stand-ins used only by the host-logic tests: a fake device population and a fake task with the same call surface
as fqlpop.population.Population / OfflineTaskWithSimulatedEvaluations. They do no FQL arithmetic -- the Trainer's
control flow (lockstep chunks, budget, pruning, score all-gather, checkpoints) is what is under test."""
import numpy as np

from fqlpop.population import INFO_KEYS
from task.task import Task


class FakePopulation:
    def __init__(self, agent_config, alphas, seeds, device="cpu"):
        self.P, self.B = len(alphas), int(agent_config.get("batch_size", 256))
        self.O = int(agent_config["ob_dims"][-1])
        self.A = int(agent_config["action_dim"])
        self.alphas = np.asarray(alphas, np.float32)
        self.seeds = list(seeds)
        self.active = list(range(self.P))
        self.steps_done = np.zeros(self.P, np.int64)
        self.calls = []
        self.device = device

    def set_active(self, members):
        self.active = [int(m) for m in members]
        self.calls.append(("active", tuple(self.active)))

    def set_dataset(self, *arrays):
        self.calls.append(("dataset", len(arrays[0])))

    def set_noise(self, **kw):
        pass

    def step(self, n=1):
        for m in self.active:
            self.steps_done[m] += n
        self.calls.append(("step", n, tuple(self.active)))

    def total_loss(self):
        self.calls.append(("total_loss", tuple(self.active)))

    def info(self, with_grad_stats=False):
        out = np.zeros((self.P, 16), np.float32)
        out[:, 0] = self.steps_done
        return out

    def info_dict(self, member, with_grad_stats=False):
        return {k: float(v) for k, v in zip(INFO_KEYS, self.info()[member])}

    def get_tree(self, member, net, which="params"):
        return {"Dense_0": {"kernel": np.full((2, 2), float(self.steps_done[member]), np.float32),
                            "bias": np.zeros(2, np.float32)}}

    def set_tree(self, member, net, tree, which="params"):
        if net == "critic" and which == "params":
            self.steps_done[member] = int(tree["Dense_0"]["kernel"][0, 0])

    def refresh_shadow(self):
        pass

    def get_counters(self):
        return self.steps_done.astype(np.int32).copy(), self.steps_done.astype(np.uint64)

    def set_counters(self, adam_steps=None, step_counter=None):
        pass

    # survivor migration: the fake member state is (steps trained, alpha, seed)
    def member_state_words(self):
        return 4

    def export_member(self, member, out=None):
        import torch
        blob = torch.tensor([float(self.steps_done[member]), float(self.alphas[member]), float(self.seeds[member]), 7.0])
        if out is not None:
            out.copy_(blob)
            return out
        return blob

    def import_member(self, member, blob):
        assert blob.numel() == 4 and float(blob[3]) == 7.0
        self.steps_done[member] = int(blob[0])
        self.alphas[member] = float(blob[1])
        self.seeds[member] = int(blob[2])
        self.calls.append(("import", member))

    def close(self):
        pass


class FakeTask(Task):
    """Scores a member by -|log10(alpha) - 2| + 1e-4 * steps trained (deterministic, alpha-ranked)."""

    def __init__(self, O=4, A=2, rows=64, device="cpu"):
        rng = np.random.default_rng(0)
        self.data = {"observations": rng.standard_normal((rows, O)).astype(np.float32),
                     "actions": rng.uniform(-1, 1, (rows, A)).astype(np.float32),
                     "next_observations": rng.standard_normal((rows, O)).astype(np.float32),
                     "rewards": -np.ones(rows, np.float32), "masks": np.ones(rows, np.float32)}
        self.device = device
        self.rollouts = 0

    def sample(self, dataset, batch_size):
        idx = np.random.randint(len(self.data["rewards"]), size=batch_size)
        return {k: v[idx] for k, v in self.data.items()}

    def device_dataset(self, split="train"):
        return [self.data[k] for k in ("observations", "actions", "next_observations", "rewards", "masks")]

    def rollout_population(self, population, seed=None, trajectory=False):
        self.rollouts += 1
        rate = np.full(population.P, np.nan, np.float32)
        for m in population.active:
            rate[m] = -abs(np.log10(population.alphas[m]) - 2.0) + 1e-4 * population.steps_done[m] / 10
        return {"success_rate": rate}

    def reset(self, seed=0):
        raise NotImplementedError

    def step(self, action):
        raise NotImplementedError

    def close(self):
        pass


class ChurnTask(FakeTask):
    """Scores that reshuffle from evaluation to evaluation (a fixed pseudo-random function of alpha and of the steps trained),
    so that SuccessiveHalving's survivors change owners often and pruned candidates re-enter (its ranking covers every
    candidate ever scored)."""

    def rollout_population(self, population, seed=None, trajectory=False):
        self.rollouts += 1
        rate = np.full(population.P, np.nan, np.float32)
        for m in population.active:
            key = int(round(float(population.alphas[m]) * 1000)) * 7919 + int(population.steps_done[m]) * 104729
            rate[m] = ((key * 2654435761) % 2 ** 32) / 2 ** 32
        return {"success_rate": rate}
