"""Parity of the CUDA population step (through the C-ABI) with the CPU oracle on identical synthetic batches,
parameters and injected noise (GPU only).

Tolerances (stated per quantity; the CUDA path multiplies in bf16 with fp32 accumulation, the reference
computes in fp32 -- SURVEY.md section 7 "bf16 vs the reference's fp32"):
  * vs the oracle with emulate_bf16=True (same operand rounding): losses/info rel 5e-3, gradients
    max-norm relative 2e-2 per leaf (tanh.approx GELU, bf16-stored act'(z)/xhat), Adam mu/nu the same,
    parameters after one step within 2.2 lr (Adam's first step is +-lr * sign(g); sign flips of ~0 grads).
  * vs the plain fp32 oracle: losses rel 3e-2, gradient cosine similarity per leaf > 0.99.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import fql_oracle as orc  # noqa: E402  (checker only)


def _make_case(P, cfg, seed=0):
    g = torch.Generator().manual_seed(seed)
    B, O, A = cfg.batch_size, cfg.ob_dim, cfg.action_dim
    batches, noises = [], []
    for _ in range(P):
        obs = torch.randn(B, O, generator=g)
        batch = {
            "observations": obs,
            "actions": (torch.rand(B, A, generator=g) * 2 - 1) * 0.99,
            "next_observations": obs + 0.05 * torch.randn(B, O, generator=g),
            "rewards": -(torch.rand(B, generator=g) < 0.9).float(),
            "masks": (torch.rand(B, generator=g) < 0.9).float(),
        }
        noise = {
            "z_critic": torch.randn(B, A, generator=g), "x0": torch.randn(B, A, generator=g),
            "t": torch.rand(B, 1, generator=g), "z_actor": torch.randn(B, A, generator=g),
            "z_mse": torch.randn(B, A, generator=g),
        }
        batches.append(batch)
        noises.append(noise)
    return batches, noises


def _np_tree(t):
    return orc.tree_map(lambda x: x.detach().cpu().numpy(), t)


def _load_population(pop, states):
    for p, st in enumerate(states):
        for net in ("critic", "actor_bc_flow", "actor_onestep_flow"):
            pop.set_tree(p, net, _np_tree(st["params"][net]), "params")
            pop.set_tree(p, net, _np_tree(st["mu"][net]), "mu")
            pop.set_tree(p, net, _np_tree(st["nu"][net]), "nu")
        pop.set_tree(p, "critic", _np_tree(st["params"]["target_critic"]), "target")
    pop.refresh_shadow()
    pop.set_counters(adam_steps=[st["step"] for st in states])


def _bind(pop, batches, noises):
    cat = lambda k: torch.cat([b[k] for b in batches], 0)
    pop.set_dataset(cat("observations"), cat("actions"), cat("next_observations"), cat("rewards"), cat("masks"))
    P, B = len(batches), batches[0]["rewards"].shape[0]
    st = lambda k: torch.stack([n[k] for n in noises], 0)
    pop.set_noise(indices=torch.arange(P * B).view(P, B), z_critic=st("z_critic"), x0=st("x0"),
                  t=st("t").squeeze(-1), z_actor=st("z_actor"), z_mse=st("z_mse"))


def _leaf_pairs(tree_a, tree_b, prefix=""):
    for k in tree_b:
        if isinstance(tree_b[k], dict):
            yield from _leaf_pairs(tree_a[k], tree_b[k], f"{prefix}/{k}")
        else:
            yield f"{prefix}/{k}", np.asarray(tree_a[k]), tree_b[k].detach().cpu().numpy()


def _cfg_dict(cfg):
    return dict(ob_dim=cfg.ob_dim, action_dim=cfg.action_dim, lr=cfg.lr, batch_size=cfg.batch_size,
                actor_hidden_dims=cfg.actor_hidden_dims, value_hidden_dims=cfg.value_hidden_dims,
                layer_norm=cfg.layer_norm, actor_layer_norm=cfg.actor_layer_norm, discount=cfg.discount,
                tau=cfg.tau, q_agg=cfg.q_agg, flow_steps=cfg.flow_steps, normalize_q_loss=cfg.normalize_q_loss,
                num_ensembles=cfg.num_ensembles)


@pytest.fixture
def chain_kernel(request, monkeypatch):
    """FQLPOP_CHAIN2 is read at graph-capture time: 0 = single-CTA layer-chain kernel, 1 = 2-CTA cluster kernel."""
    monkeypatch.setenv("FQLPOP_CHAIN2", str(request.param))
    return request.param


@pytest.mark.parametrize("chain_kernel", [0, 1], indirect=True, ids=["single_cta", "cluster"])
@pytest.mark.parametrize("shape,variant", [((28, 5), "default"), ((42, 8), "min_norm"), ((37, 5), "actor_ln")])
def test_one_update_matches_oracle(shape, variant, chain_kernel):
    from fqlpop.population import INFO_KEYS, Population
    O, A = shape
    cfg = orc.FQLConfig(ob_dim=O, action_dim=A, batch_size=256)
    if variant == "min_norm":
        cfg.q_agg, cfg.normalize_q_loss, cfg.discount = "min", True, 0.995
    if variant == "actor_ln":
        cfg.actor_layer_norm = True
    alphas = [3.0, 30.0, 300.0]
    seeds = [11, 22, 33]
    P = len(alphas)
    states = [orc.init_state(s, cfg) for s in seeds]
    # non-trivial biases / LayerNorm parameters and a target that differs from the critic
    g = torch.Generator().manual_seed(5)
    for st in states:
        for net in st["params"]:
            for name, leaf in orc.tree_leaves(st["params"][net]):
                if name.endswith("bias"):
                    leaf.add_(0.05 * torch.randn(leaf.shape, generator=g))
                if name.endswith("scale"):
                    leaf.add_(0.1 * torch.randn(leaf.shape, generator=g))
        for name, leaf in orc.tree_leaves(st["params"]["target_critic"]):
            leaf.add_(0.01 * torch.randn(leaf.shape, generator=g))
    batches, noises = _make_case(P, cfg)
    pop = Population(_cfg_dict(cfg), alphas, seeds)
    _load_population(pop, states)
    _bind(pop, batches, noises)
    pop.step(1)
    info = pop.info(with_grad_stats=True)

    for p in range(P):
        new_e, info_e, grads_e = orc.update(states[p], batches[p], noises[p], cfg, alphas[p], emulate_bf16=True,
                                            return_grads=True)
        _, info_f, grads_f = orc.update(states[p], batches[p], noises[p], cfg, alphas[p], emulate_bf16=False,
                                        return_grads=True)
        for i, k in enumerate(INFO_KEYS):
            got, want, want_f = float(info[p, i]), info_e[k], info_f[k]
            tol = 5e-3 * max(1.0, abs(want)) if not k.startswith("grad/") else 3e-2 * max(1e-3, abs(want))
            assert abs(got - want) <= tol, f"member {p} {k}: cuda {got} oracle(bf16) {want}"
            if not k.startswith("grad/"):
                assert abs(got - want_f) <= 3e-2 * max(1.0, abs(want_f)), f"member {p} {k}: cuda {got} oracle(fp32) {want_f}"
        for net in ("critic", "actor_bc_flow", "actor_onestep_flow"):
            g_c = pop.get_tree(p, net, "grad")
            for name, a, b in _leaf_pairs(g_c, grads_e[net], net):
                scale = np.abs(b).max() + 1e-12
                err = np.abs(a - b).max() / scale
                assert err < 2e-2, f"member {p} grad {name}: max-norm rel err {err:.3e} (scale {scale:.3e})"
            for name, a, b in _leaf_pairs(g_c, grads_f[net], net):
                cos = float((a * b).sum() / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-30))
                assert cos > 0.99, f"member {p} grad {name}: cosine vs fp32 oracle {cos:.4f}"
            for which, key in (("mu", "mu"), ("nu", "nu")):
                t_c = pop.get_tree(p, net, which)
                for name, a, b in _leaf_pairs(t_c, new_e[key][net], net):
                    scale = np.abs(b).max() + 1e-20
                    assert np.abs(a - b).max() / scale < 4e-2, f"member {p} {which} {name}"
            t_c = pop.get_tree(p, net, "params")
            for name, a, b in _leaf_pairs(t_c, new_e["params"][net], net):
                assert np.abs(a - b).max() <= 2.2 * cfg.lr, f"member {p} params {name}: {np.abs(a - b).max():.3e}"
                # the bulk of the entries moves exactly like the oracle
                frac = float((np.abs(a - b) < 0.2 * cfg.lr).mean())
                assert frac > 0.97, f"member {p} params {name}: only {frac:.3f} within 0.2 lr"
        t_c = pop.get_tree(p, "critic", "target")
        for name, a, b in _leaf_pairs(t_c, new_e["params"]["target_critic"], "target"):
            assert np.abs(a - b).max() <= 2.2 * cfg.lr * cfg.tau + 1e-7, f"member {p} target {name}"
    steps, ctr = pop.get_counters()
    assert list(steps) == [1] * P and list(ctr) == [1] * P
    pop.close()


def test_total_loss_matches_update_info_and_leaves_state_untouched():
    from fqlpop.population import INFO_KEYS, Population
    cfg = orc.FQLConfig(ob_dim=28, action_dim=5, batch_size=128)
    alphas, seeds = [10.0, 100.0], [1, 2]
    states = [orc.init_state(s, cfg) for s in seeds]
    batches, noises = _make_case(2, cfg, seed=3)
    pop = Population(_cfg_dict(cfg), alphas, seeds)
    _load_population(pop, states)
    _bind(pop, batches, noises)
    before = pop.get_tree(0, "critic", "params")
    pop.total_loss()
    info = pop.info()
    after = pop.get_tree(0, "critic", "params")
    for k in before:
        for kk in before[k]:
            assert np.array_equal(before[k][kk], after[k][kk])
    for p in range(2):
        with torch.no_grad():
            pass
        params = orc.tree_map(lambda x: x.clone(), states[p]["params"])
        _, info_e = orc.total_loss(params, batches[p], noises[p], cfg, alphas[p], emulate_bf16=True)
        for i, k in enumerate(INFO_KEYS[:10]):
            want = float(info_e[k])
            assert abs(float(info[p, i]) - want) <= 5e-3 * max(1.0, abs(want)), (p, k, float(info[p, i]), want)
    steps, ctr = pop.get_counters()
    assert list(steps) == [0, 0] and list(ctr) == [0, 0]
    pop.close()


@pytest.mark.gpu
def test_cluster_and_single_cta_kernels_agree_over_steps(monkeypatch):
    """Same population, same explicit batch and noise, 5 updates with each layer-chain kernel: the two compute the same
    arithmetic (different accumulation order over K only), so info and parameters stay close."""
    from fqlpop.population import Population
    cfg = orc.FQLConfig(ob_dim=28, action_dim=5, batch_size=256)
    alphas, seeds = [3.0, 30.0, 300.0, 1000.0], [1, 2, 3, 4]
    batches, noises = _make_case(len(alphas), cfg)
    states = [orc.init_state(s, cfg) for s in seeds]
    res = []
    for mode in ("0", "1"):
        monkeypatch.setenv("FQLPOP_CHAIN2", mode)
        pop = Population(_cfg_dict(cfg), alphas, seeds)
        _load_population(pop, states)
        _bind(pop, batches, noises)
        pop.step(5)
        res.append((pop.info().copy(), [pop.get_tree(p, "critic", "params") for p in range(len(alphas))],
                    [pop.get_tree(p, "actor_onestep_flow", "params") for p in range(len(alphas))]))
        pop.close()
    (i0, c0, a0), (i1, c1, a1) = res
    # 5 steps amplify the bf16 rounding differences of the two accumulation orders: 1e-2 (the one-step test holds both
    # kernels to 5e-3 against the oracle)
    assert np.allclose(i0[:, :10], i1[:, :10], rtol=1e-2, atol=1e-2), np.abs(i0[:, :10] - i1[:, :10]).max()
    for t0, t1 in zip(c0 + a0, c1 + a1):
        for name, x, y in _leaf_pairs(t0, {k: {kk: torch.as_tensor(vv) for kk, vv in v.items()} for k, v in t1.items()}):
            assert np.abs(x - y).max() <= 5 * 2.2 * cfg.lr, name


@pytest.mark.parametrize("B,P,active", [(192, 2, None), (1024, 2, None), (64, 3, None), (256, 6, [0, 2, 5]), (256, 5, None)],
                         ids=["ragged_B192", "cube_double_B1024", "small_B64", "active_subset", "ragged_group_P5"])
def test_update_edge_shapes_match_oracle(B, P, active):
    """Batch sizes that do not fill a 128-row tile (192, 64), the B=1024 configuration of BASELINE config 5, a population
    that does not fill the last member group (5 with groups of 4), and a pruned population (successive halving leaves
    members inactive): losses / info of every ACTIVE member against the oracle, parameters inside the Adam bound, and
    every bit of an INACTIVE member's state untouched."""
    from fqlpop.population import INFO_KEYS, Population
    O, A = (37, 5) if B == 1024 else (28, 5)
    cfg = orc.FQLConfig(ob_dim=O, action_dim=A, batch_size=B)
    alphas = list(np.logspace(np.log10(3), np.log10(1000), P))
    seeds = list(range(40, 40 + P))
    states = [orc.init_state(s, cfg) for s in seeds]
    g = torch.Generator().manual_seed(6)
    for st in states:
        for net in st["params"]:
            for name, leaf in orc.tree_leaves(st["params"][net]):
                if name.endswith("bias"):
                    leaf.add_(0.05 * torch.randn(leaf.shape, generator=g))
    batches, noises = _make_case(P, cfg, seed=3)
    pop = Population(_cfg_dict(cfg), alphas, seeds)
    _load_population(pop, states)
    _bind(pop, batches, noises)
    members = list(range(P)) if active is None else active
    if active is not None:
        pop.set_active(active)
    before = {p: {net: pop.get_tree(p, net, "params") for net in ("critic", "actor_bc_flow", "actor_onestep_flow")}
              for p in range(P) if p not in members}
    pop.step(1)
    info = pop.info(with_grad_stats=False)
    for p in members:
        new_e, info_e, _ = orc.update(states[p], batches[p], noises[p], cfg, alphas[p], emulate_bf16=True, return_grads=True)
        for i, k in enumerate(INFO_KEYS):
            if k.startswith("grad/"):
                continue
            got, want = float(info[p, i]), info_e[k]
            assert abs(got - want) <= 5e-3 * max(1.0, abs(want)), f"B={B} member {p} {k}: cuda {got} oracle {want}"
        for net in ("critic", "actor_bc_flow", "actor_onestep_flow"):
            for name, a, b in _leaf_pairs(pop.get_tree(p, net, "params"), new_e["params"][net], net):
                assert np.abs(a - b).max() <= 2.2 * cfg.lr, f"B={B} member {p} params {name}"
                assert float((np.abs(a - b) < 0.2 * cfg.lr).mean()) > 0.97, f"B={B} member {p} params {name}"
    for p, trees in before.items():
        for net, tree in trees.items():
            after = pop.get_tree(p, net, "params")
            for layer in tree:
                for leaf in tree[layer]:
                    assert np.array_equal(tree[layer][leaf], after[layer][leaf]), f"inactive member {p} {net}/{layer}/{leaf} changed"
    steps, _ = pop.get_counters()
    assert list(steps) == [1 if p in members else 0 for p in range(P)]
    pop.close()


def test_batch_size_must_be_a_multiple_of_64():
    """Documented limit of the weight-gradient kernels; the reference's batch sizes (256, 1024) satisfy it. It must fail
    loudly at create time, not silently pad."""
    from fqlpop._lib import FqlPopError
    from fqlpop.population import Population
    cfg = orc.FQLConfig(ob_dim=28, action_dim=5, batch_size=200)
    with pytest.raises(FqlPopError, match="multiple of 64"):
        Population(_cfg_dict(cfg), [3.0], [0])
