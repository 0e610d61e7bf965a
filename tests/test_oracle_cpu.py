"""Pins of the CPU oracle that need no GPU: committed regression fixtures of its own outputs (tests/golden/
oracle_small.npz -- oracle-generated, NOT reference-generated: the FQL submodule is absent, see oracle header), and
independent derivations of its pieces (torch.optim.Adam, finite differences, a literal host loop of the reference's
evaluate_actor_fn / Task.step)."""
import os

import numpy as np
import pytest
import torch

from oracle import fql_oracle as orc

GOLD = os.path.join(os.path.dirname(__file__), "golden", "oracle_small.npz")


def small_case(dtype=torch.float64):
    cfg = orc.FQLConfig(ob_dim=7, action_dim=3, batch_size=16, actor_hidden_dims=(32, 32), value_hidden_dims=(32, 32),
                        flow_steps=4)
    g = torch.Generator().manual_seed(123)
    B, O, A = cfg.batch_size, cfg.ob_dim, cfg.action_dim
    obs = torch.randn(B, O, generator=g, dtype=dtype)
    batch = {"observations": obs, "actions": torch.rand(B, A, generator=g, dtype=dtype) * 2 - 1,
             "next_observations": obs + 0.05 * torch.randn(B, O, generator=g, dtype=dtype),
             "rewards": -(torch.rand(B, generator=g, dtype=dtype) > 0.2).to(dtype), "masks": torch.ones(B, dtype=dtype)}
    batch["masks"] = (batch["rewards"] != 0).to(dtype)
    noise = {k: torch.randn(B, A, generator=g, dtype=dtype) for k in ("z_critic", "x0", "z_actor", "z_mse")}
    noise["t"] = torch.rand(B, 1, generator=g, dtype=dtype)
    return cfg, orc.init_state(5, cfg, dtype), batch, noise


def test_oracle_regression_fixture():
    cfg, st, batch, noise = small_case()
    new, info = orc.update(st, batch, noise, cfg, alpha=30.0)
    new2, info2 = orc.update(new, batch, noise, cfg, alpha=30.0)
    got = {f"info1/{k}": v for k, v in info.items()} | {f"info2/{k}": v for k, v in info2.items()}
    got["w_critic"] = new2["params"]["critic"]["Dense_1"]["kernel"].numpy()
    got["w_target"] = new2["params"]["target_critic"]["Dense_0"]["kernel"].numpy()
    got["w_pi"] = new2["params"]["actor_onestep_flow"]["Dense_2"]["kernel"].numpy()
    if os.environ.get("FQLPOP_WRITE_GOLDEN"):
        np.savez(GOLD, **{k: np.asarray(v) for k, v in got.items()})
    want = np.load(GOLD)
    for k in want.files:
        np.testing.assert_allclose(np.asarray(got[k]), want[k], rtol=1e-9, atol=1e-12, err_msg=k)


def test_update_is_adam_plus_polyak_of_the_loss_gradient():
    cfg, st, batch, noise = small_case()
    names = ("critic", "actor_bc_flow", "actor_onestep_flow")
    work = {k: orc.tree_map(lambda x: x.clone().requires_grad_(k in names), st["params"][k]) for k in st["params"]}
    leaves = [t for k in names for _, t in orc.tree_leaves(work[k])]
    opt = torch.optim.Adam(leaves, lr=cfg.lr, betas=(cfg.adam_b1, cfg.adam_b2), eps=cfg.adam_eps)
    state = st
    for _ in range(3):
        opt.zero_grad()
        loss, _ = orc.total_loss(work, batch, noise, cfg, alpha=7.0)
        loss.backward()
        opt.step()
        with torch.no_grad():  # Polyak with the post-Adam critic
            for (_, t), (_, c) in zip(orc.tree_leaves(work["target_critic"]), orc.tree_leaves(work["critic"])):
                t.mul_(1 - cfg.tau).add_(cfg.tau * c)
        state, _ = orc.update(state, batch, noise, cfg, alpha=7.0)
    for k in names:
        for (n, a), (_, b) in zip(orc.tree_leaves(work[k]), orc.tree_leaves(state["params"][k])):
            # torch.optim.Adam puts eps outside the bias-corrected sqrt exactly like optax.adam
            np.testing.assert_allclose(a.detach().numpy(), b.numpy(), rtol=1e-7, atol=1e-10, err_msg=f"{k}/{n}")
    for (n, a), (_, b) in zip(orc.tree_leaves(work["target_critic"]), orc.tree_leaves(state["params"]["target_critic"])):
        np.testing.assert_allclose(a.detach().numpy(), b.numpy(), rtol=1e-7, atol=1e-10, err_msg=f"target/{n}")


def test_loss_terms_by_hand():
    cfg, st, batch, noise = small_case()
    p = st["params"]
    _, info = orc.total_loss(p, batch, noise, cfg, alpha=10.0)
    obs, act = batch["observations"], batch["actions"]
    # critic: y = r + gamma * mask * mean_e Qbar(s', clip(pi(s', z)))
    a2 = orc.sample_actions(p, batch["next_observations"], noise["z_critic"], cfg)
    assert float(a2.abs().max()) <= 1.0
    y = batch["rewards"] + cfg.discount * batch["masks"] * orc.critic_apply(p["target_critic"], batch["next_observations"], a2, cfg).mean(0)
    q = orc.critic_apply(p["critic"], obs, act, cfg)
    assert q.shape == (cfg.num_ensembles, cfg.batch_size)
    assert abs(float(((q - y) ** 2).mean()) - float(info["critic/critic_loss"])) < 1e-12
    # Euler flow: x <- x + v(s, x, i/F)/F, F steps, clipped
    x = noise["z_actor"].clone()
    for i in range(cfg.flow_steps):
        t = torch.full((cfg.batch_size, 1), i / cfg.flow_steps, dtype=x.dtype)
        x = x + orc.actor_apply(p["actor_bc_flow"], obs, x, t, cfg) / cfg.flow_steps
    tgt = x.clamp(-1, 1)
    np.testing.assert_allclose(orc.compute_flow_actions(p, obs, noise["z_actor"], cfg).numpy(), tgt.numpy(), rtol=1e-12)
    a_pi = orc.actor_apply(p["actor_onestep_flow"], obs, noise["z_actor"], None, cfg)
    assert abs(float(((a_pi - tgt) ** 2).mean()) - float(info["actor/distill_loss"])) < 1e-12
    total = info["actor/bc_flow_loss"] + 10.0 * info["actor/distill_loss"] + info["actor/q_loss"]
    assert abs(float(total) - float(info["actor/actor_loss"])) < 1e-12


def test_q_agg_min_and_normalize_q_loss():
    cfg, st, batch, noise = small_case()
    cfg.q_agg, cfg.normalize_q_loss = "min", True
    _, info = orc.total_loss(st["params"], batch, noise, cfg)
    p = st["params"]
    qa = orc.critic_apply(p["critic"], batch["observations"],
                          orc.actor_apply(p["actor_onestep_flow"], batch["observations"], noise["z_actor"], None, cfg).clamp(-1, 1), cfg).mean(0)
    assert abs(float(-qa.mean() / qa.abs().mean()) - float(info["actor/q_loss"])) < 1e-12


class _HostSimTask:
    """Literal restatement of OfflineTaskWithSimulatedEvaluations.reset/step
    (/root/reference/task/offline_task_simulated.py:60-107) on top of the oracle's network functions."""

    def __init__(self, env, init_obs, max_episode_steps):
        self.env, self.init, self.max_episode_steps = env, init_obs, max_episode_steps

    def reset(self, seed=None):
        self.cur = self.init.clone()
        self.episode_steps = 0
        self.invalidate = [False] * len(self.init)
        return self.cur.numpy(), [{} for _ in self.invalidate]

    def step(self, actions):
        a = torch.as_tensor(actions, dtype=self.cur.dtype)
        nxt = orc.state_predictor_apply(self.env["state"], self.cur, a)
        term = (orc.termination_predictor_apply(self.env["termination"], nxt) > 0).numpy()
        self.cur = nxt
        self.episode_steps += 1
        trunc = np.full_like(term, self.episode_steps >= self.max_episode_steps, dtype=bool)
        infos = [{"success": bool(s)} for s in term]
        for i in range(len(infos)):
            if self.invalidate[i]:
                infos[i]["invalid"] = True
            if term[i] or trunc[i]:
                self.invalidate[i] = True
        return nxt.numpy(), np.where(term, 0, -1), term, trunc, infos


@pytest.mark.parametrize("max_steps", [3, 40])
def test_oracle_rollout_equals_the_reference_host_loop(max_steps):
    from evaluator.evaluation import evaluate_actor_fn
    cfg = orc.FQLConfig(ob_dim=6, action_dim=2, actor_hidden_dims=(32, 32), value_hidden_dims=(32, 32))
    params = orc.init_params(3, cfg, torch.float64)
    env = orc.init_envmodel(4, 6, 2, hidden=(16, 16), dtype=torch.float64)
    env["termination"]["Dense_2"]["bias"] -= 1.0   # episodes end at different steps; with max_steps=3 some only truncate
    g = torch.Generator().manual_seed(0)
    n = 24
    init = torch.randn(n, 6, generator=g, dtype=torch.float64)
    noise = torch.randn(max_steps, n, 2, generator=g, dtype=torch.float64)
    want = orc.rollout(params, env, init, noise, cfg, max_steps)
    step = {"i": 0}

    def actor(observations, temperature):
        z = noise[step["i"]]
        step["i"] += 1
        return orc.sample_actions(params, torch.as_tensor(observations), z, cfg).numpy()

    stats, transitions = evaluate_actor_fn(actor, _HostSimTask(env, init, max_steps))
    assert len(transitions) == want["steps"]
    assert abs(float(stats["success"]) - want["success_rate"]) < 1e-6  # the oracle reports a float32 mean
    assert (0 < want["success_rate"] < 1) if max_steps == 3 else (want["success_rate"] == 1.0 and want["steps"] > 3)
