"""Parity on BASELINE.json's own configurations (GPU only) -- the shapes the benchmark runs, with the DEFAULT kernel
selection (no FQLPOP_* switch set):

  * config 2: lockstep population of 16 alpha values, cube-single shape, batch 256 (four graph groups of 4) -- one update
    of all 16 members vs the oracle; also 8 and 9 members (either side of the population size at which the default
    switches from the first-generation N-split cluster kernel to the second-generation pair kernel) and 12;
  * per-step losses over 200 updates of two members vs the fp32 oracle fed the same batches and noise (north_star:
    "per-step losses ... within a stated fp tolerance"): the envelope is written below;
  * config 3: 4096 episodes x 200 steps, antsoccer shape, termination bias calibrated so that about half of the episodes
    terminate, success rate / mean return / executed steps vs the oracle's rollout for two members.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import fql_oracle as orc  # noqa: E402  (checker only)
from test_update_gpu import _bind, _cfg_dict, _leaf_pairs, _load_population, _make_case  # noqa: E402


@pytest.fixture(autouse=True)
def _default_kernel_selection(monkeypatch):
    for k in ("FQLPOP_CHAIN2", "FQLPOP_CHAIN_GEN", "FQLPOP_CHAIN_PAIR", "FQLPOP_GROUP_SIZE", "FQLPOP_ADAM_CTAS"):
        monkeypatch.delenv(k, raising=False)


# B (a multiple of 64): 256 is the benchmark's batch; 192 gives the pair kernel a half-empty second tile (its TMA stores
# must be clipped by the tensor map), 320 an odd number of 128-row tiles (one CTA per tile then) with a ragged last one;
# 1024 is config 5's batch.
@pytest.mark.parametrize("P,fuse,B", [(16, 0, 256), (8, 0, 256), (9, 0, 256), (12, 0, 256), (5, 1, 256),
                                      (12, 0, 192), (10, 0, 320), (10, 0, 1024)])
def test_population_update_default_kernels_matches_oracle(P, fuse, B, monkeypatch):
    from fqlpop.population import INFO_KEYS, Population
    monkeypatch.setenv("FQLPOP_FUSE_ADAM", str(fuse))   # 1: Adam + Polyak inside the weight-gradient GEMM's epilogue
    cfg = orc.FQLConfig(ob_dim=28, action_dim=5, batch_size=B)
    alphas = list(np.logspace(np.log10(3), np.log10(1000), P))
    seeds = list(range(100, 100 + P))
    states = [orc.init_state(s, cfg) for s in seeds]
    g = torch.Generator().manual_seed(8)
    for st in states:
        for net in st["params"]:
            for name, leaf in orc.tree_leaves(st["params"][net]):
                if name.endswith("bias"):
                    leaf.add_(0.05 * torch.randn(leaf.shape, generator=g))
                if name.endswith("scale"):
                    leaf.add_(0.1 * torch.randn(leaf.shape, generator=g))
    batches, noises = _make_case(P, cfg, seed=12)
    pop = Population(_cfg_dict(cfg), alphas, seeds)
    _load_population(pop, states)
    _bind(pop, batches, noises)
    pop.step(1)
    info = pop.info(with_grad_stats=False)
    for p in range(P):
        new_e, info_e = orc.update(states[p], batches[p], noises[p], cfg, alphas[p], emulate_bf16=True)
        for i, k in enumerate(INFO_KEYS):
            if k.startswith("grad/"):
                continue
            got, want = float(info[p, i]), info_e[k]
            assert abs(got - want) <= 5e-3 * max(1.0, abs(want)), f"P={P} member {p} {k}: cuda {got} oracle {want}"
        if p in (0, P // 2, P - 1):   # parameters of a spread of members (reading all 16 back is slow, not more telling)
            for net in ("critic", "actor_bc_flow", "actor_onestep_flow"):
                for name, a, b in _leaf_pairs(pop.get_tree(p, net, "params"), new_e["params"][net], net):
                    assert np.abs(a - b).max() <= 2.2 * cfg.lr, f"P={P} member {p} params {name}"
                    assert float((np.abs(a - b) < 0.2 * cfg.lr).mean()) > 0.97, f"P={P} member {p} params {name}"
    pop.close()


def test_two_hundred_updates_track_the_fp32_oracle():
    """Per-step losses of the bf16 CUDA path against the fp32 oracle over 200 updates of two members, every step fed the
    same batch and noise in both. Envelope (stated, checked at every step): critic_loss, bc_flow_loss, distill_loss and
    q_loss within 5 % relative (+ 5e-3 absolute) of the oracle's value at that step; at the end the two loss curves'
    means over the last 50 steps agree within 2 %."""
    from fqlpop.population import INFO_KEYS, Population
    cfg = orc.FQLConfig(ob_dim=28, action_dim=5, batch_size=256)
    alphas, seeds = [10.0, 300.0], [7, 8]
    P, steps = 2, 200
    states = [orc.init_state(s, cfg) for s in seeds]
    pop = Population(_cfg_dict(cfg), alphas, seeds)
    _load_population(pop, states)
    # a fixed pool of 4096 transitions per member; every step draws its batch rows and noise from a seeded generator
    g = torch.Generator().manual_seed(21)
    N, B, O, A = 4096, cfg.batch_size, cfg.ob_dim, cfg.action_dim
    obs = torch.randn(P * N, O, generator=g)
    data = {"observations": obs, "actions": (torch.rand(P * N, A, generator=g) * 2 - 1) * 0.99,
            "next_observations": obs + 0.05 * torch.randn(P * N, O, generator=g),
            "rewards": -(torch.rand(P * N, generator=g) < 0.9).float(),
            "masks": (torch.rand(P * N, generator=g) < 0.9).float()}
    pop.set_dataset(*(data[k] for k in ("observations", "actions", "next_observations", "rewards", "masks")))
    keys = ("critic/critic_loss", "actor/bc_flow_loss", "actor/distill_loss", "actor/q_loss")
    idx_of = {k: INFO_KEYS.index(k) for k in keys}
    got = np.zeros((steps, P, len(keys)))
    want = np.zeros((steps, P, len(keys)))
    for s in range(steps):
        idx = torch.stack([p * N + torch.randint(0, N, (B,), generator=g) for p in range(P)])
        noise = {k: torch.randn(P, B, A, generator=g) for k in ("z_critic", "x0", "z_actor", "z_mse")}
        noise["t"] = torch.rand(P, B, generator=g)
        pop.set_noise(indices=idx, **noise)
        pop.step(1)
        info = pop.info()
        for p in range(P):
            batch = {k: v[idx[p]] for k, v in data.items()}
            nz = {k: (v[p] if k != "t" else v[p].unsqueeze(-1)) for k, v in noise.items()}
            states[p], info_e = orc.update(states[p], batch, nz, cfg, alphas[p], emulate_bf16=False)
            for j, k in enumerate(keys):
                got[s, p, j] = float(info[p, idx_of[k]])
                want[s, p, j] = float(info_e[k])
    rel = np.abs(got - want) / (np.abs(want) + 0.1)   # 0.1: losses near zero are compared absolutely
    worst = np.unravel_index(np.argmax(rel), rel.shape)
    assert rel.max() < 0.05, f"step {worst[0]} member {worst[1]} {keys[worst[2]]}: cuda {got[worst]} oracle {want[worst]}"
    tail_g, tail_w = got[-50:].mean(0), want[-50:].mean(0)
    assert np.all(np.abs(tail_g - tail_w) <= 0.02 * np.abs(tail_w) + 2e-3), (tail_g, tail_w)
    assert np.isfinite(got).all()
    pop.close()


def test_config3_rollout_at_scale_matches_oracle():
    """4096 episodes x 200 steps, antsoccer shape, ~50 % of the episodes terminating: the fused kernel's success rate, mean
    return and executed lockstep steps vs the oracle's restatement of evaluate_actor_fn + Task.step on the SAME initial
    observations and policy noise. Trajectories of a random 200-step dynamical system are not comparable step by step in
    bf16 vs fp32 (the T=12 test does that); the aggregate statistics are: success rate within 0.03 absolute, mean
    return within 3 % -- about 3 sigma of the sampling noise of 4096 Bernoulli episodes (0.5 / sqrt(4096) = 0.008)."""
    from fqlpop.envmodel import EnvModel
    from fqlpop.population import Population
    from test_rollout_gpu import _envmodel, _np_tree, _population
    O, A, n, T = 42, 8, 4096, 200
    cfg = orc.FQLConfig(ob_dim=O, action_dim=A, batch_size=256)
    pop, states = _population(cfg, [10.0, 100.0], [3, 4])
    g = torch.Generator().manual_seed(17)
    init = torch.randn(n, O, generator=g)
    noise = torch.randn(2, T, n, A, generator=g)
    # termination bias by bisection on the ORACLE (member 0, 512 episodes) so that about half terminate before T
    lo, hi = -20.0, 20.0
    for _ in range(14):
        mid = 0.5 * (lo + hi)
        em = _envmodel(O, A, term_bias=mid)
        r = orc.rollout(states[0]["params"], em, init[:512], noise[0, :, :512], cfg, T)["success_rate"]
        if abs(r - 0.5) < 0.05:
            break
        lo, hi = (lo, mid) if r > 0.5 else (mid, hi)
    assert 0.3 < r < 0.7, f"calibration failed: {r}"
    dem = EnvModel(_np_tree(em["state"]), _np_tree(em["termination"]), O, A)
    out = pop.rollout(dem, init, max_episode_steps=T, noise=noise)
    torch.cuda.synchronize()
    for p in range(2):
        ref = orc.rollout(states[p]["params"], em, init, noise[p], cfg, T, emulate_bf16=True)
        sr, want_sr = float(out["success_rate"][p]), ref["success_rate"]
        assert 0.2 < want_sr < 0.8
        assert abs(sr - want_sr) <= 0.03, f"member {p}: success rate cuda {sr} oracle {want_sr}"
        ret, want_ret = float(out["return_mean"][p]), float(ref["returns"].mean())
        assert abs(ret - want_ret) <= 0.03 * abs(want_ret), f"member {p}: mean return cuda {ret} oracle {want_ret}"
        assert ref["steps"] <= int(out["steps"][p]) <= min(T, ref["steps"] + 1)
        # per-episode outcomes agree for the large majority (bf16 rounding flips decisions that sit on the threshold)
        agree = float((out["success"][p].cpu().bool() == ref["success"]).float().mean())
        assert agree > 0.9, f"member {p}: only {agree:.3f} of the episodes end the same way"
    pop.close()
    dem.close()


@pytest.mark.parametrize("fuse", [0, 1], ids=["adam_kernel", "adam_in_wgrad_epilogue"])
def test_multi_step_call_equals_single_step_calls(fuse, monkeypatch):
    """A call of n steps must leave exactly the state n single-step calls leave, and the gradients of the LAST step must be
    readable afterwards -- with the optimizer as its own kernel (default) and fused into the weight-gradient GEMM's
    epilogue (FQLPOP_FUSE_ADAM=1: the first n - 1 steps of a call then replay a graph variant that does not store the
    weight gradients)."""
    from fqlpop.population import Population
    monkeypatch.setenv("FQLPOP_FUSE_ADAM", str(fuse))
    cfg = orc.FQLConfig(ob_dim=28, action_dim=5, batch_size=256)
    alphas, seeds = [3.0, 30.0, 300.0, 1000.0], [1, 2, 3, 4]
    g = torch.Generator().manual_seed(2)
    N, O, A = 8192, cfg.ob_dim, cfg.action_dim
    obs = torch.randn(N, O, generator=g)
    data = (obs, (torch.rand(N, A, generator=g) * 2 - 1) * 0.99, obs + 0.05 * torch.randn(N, O, generator=g),
            -(torch.rand(N, generator=g) < 0.9).float(), (torch.rand(N, generator=g) < 0.9).float())
    pops = []
    for _ in range(2):
        pop = Population(_cfg_dict(cfg), alphas, seeds)
        pop.set_dataset(*data)
        pop.set_noise()           # device Philox: rows and noise are a function of (seed, step counter)
        pops.append(pop)
    pops[0].step(6)
    for _ in range(6):
        pops[1].step(1)
    torch.cuda.synchronize()
    for p in range(len(alphas)):
        for net in ("critic", "actor_bc_flow", "actor_onestep_flow"):
            for which in ("params", "mu", "nu", "grad") + (("target",) if net == "critic" else ()):
                ta, tb = pops[0].get_tree(p, net, which), pops[1].get_tree(p, net, which)
                for layer in ta:
                    for leaf in ta[layer]:
                        np.testing.assert_array_equal(ta[layer][leaf], tb[layer][leaf], err_msg=f"{p} {net} {which} {layer}/{leaf}")
    np.testing.assert_allclose(pops[0].info()[:, :10], pops[1].info()[:, :10], rtol=1e-5, atol=1e-6)   # float atomics
    for pop in pops:
        pop.close()
