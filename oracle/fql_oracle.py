"""CPU oracle for the population FQL hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module, and only as the checker / the timed CPU baseline. The product path (flow-q-learning_b200/) never
imports it and has no CPU fallback.

What it restates, and where it comes from
-----------------------------------------
* FQL agent update (critic TD loss with target network, flow-matching BC loss, F-step Euler flow feeding
  the alpha-weighted one-step distillation loss, -Q actor term, Adam, Polyak target update).
  The arithmetic lives in the git submodule `fql/` (MazenAmria/fql @ b77fec4e69e7ca81519b5d7e098c8b73d480db35,
  /root/reference/.gitmodules:1-4, /root/reference/.SUBMODULES.json:8-15) which is ABSENT from
  /root/reference, and jax/flax/optax are not installable here. The algorithm below is therefore a
  restatement of the published upstream FQL implementation (seohongpark/fql: agents/fql.py,
  utils/networks.py, utils/flax_utils.py; paper arXiv 2502.02538 cited at /root/reference/README.md:3),
  anchored on the reference's own call sites and pinned hyper-parameters:
    - FQLAgent.create(seed, ex_obs, ex_act, config)      /root/reference/trainer/experiment.py:44-49
    - agent.update(batch) -> (agent, info)               /root/reference/trainer/experiment.py:109
    - agent.total_loss(batch, grad_params=None)          /root/reference/trainer/experiment.py:115
    - agent.sample_actions(observations, seed, temp)     /root/reference/evaluator/evaluation.py:58-63,94
    - lr, batch_size, hidden dims, layer_norm, discount, tau, q_agg, alpha, flow_steps,
      normalize_q_loss defaults                          /root/reference/trainer/config.py:6-23
  PARITY UNPINNED for this part: the reference holds no tests, golden vectors or checkpoints for it
  (SURVEY.md section 0.5), and the reference itself cannot be executed here.
* World-model transition + termination + evaluation bookkeeping (fully specified by reference sources):
    - BaselineStatePredictor.__call__                    /root/reference/envmodel/baseline.py:25-37
    - TerminationPredictor.__call__                      /root/reference/envmodel/termination_predictor.py:14-21
    - OfflineTaskWithSimulatedEvaluations.step           /root/reference/task/offline_task_simulated.py:85-107
    - evaluate_actor_fn                                  /root/reference/evaluator/evaluation.py:67-114
  Also unpinned by execution (flax is absent) but there is no recalled content: every line follows the
  files above.

Every random draw is an explicit argument (JAX threefry streams cannot be reproduced here), so the CUDA
path and this oracle consume identical noise.

`emulate_bf16=True` rounds every GEMM operand (activations, weights, back-propagated gradients) to
bfloat16 exactly where the CUDA path does, while accumulating in the working dtype; it lets the parity
tests separate precision effects from logic errors.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, Optional, Tuple

import numpy as np
import torch

NETS = ("critic", "target_critic", "actor_bc_flow", "actor_onestep_flow")
INFO_KEYS = (
    "critic/critic_loss", "critic/q_mean", "critic/q_max", "critic/q_min",
    "actor/actor_loss", "actor/bc_flow_loss", "actor/distill_loss", "actor/q_loss", "actor/q", "actor/mse",
    "grad/max", "grad/min", "grad/norm",
)


@dataclass
class FQLConfig:
    """Mirror of AgentConfig (/root/reference/trainer/config.py:6-23)."""
    ob_dim: int = 28
    action_dim: int = 5
    lr: float = 3e-4
    batch_size: int = 256
    actor_hidden_dims: Tuple[int, ...] = (512, 512, 512, 512)
    value_hidden_dims: Tuple[int, ...] = (512, 512, 512, 512)
    layer_norm: bool = True
    actor_layer_norm: bool = False
    discount: float = 0.99
    tau: float = 0.005
    q_agg: str = "mean"
    alpha: float = 10.0
    flow_steps: int = 10
    normalize_q_loss: bool = False
    num_ensembles: int = 2
    ln_eps: float = 1e-6
    adam_b1: float = 0.9
    adam_b2: float = 0.999
    adam_eps: float = 1e-8


# --------------------------------------------------------------------------------------------------
# parameters
# --------------------------------------------------------------------------------------------------
def _glorot_uniform(gen: torch.Generator, fan_in: int, fan_out: int, shape, dtype):
    # flax variance_scaling(1.0, 'fan_avg', 'uniform'): U(-l, l), l = sqrt(3 / fan_avg)
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return ((torch.rand(shape, generator=gen, dtype=torch.float64) * 2 - 1) * lim).to(dtype)


def _init_mlp(gen, in_dim, hidden, out_dim, layer_norm, ens, dtype):
    p: Dict[str, Dict[str, torch.Tensor]] = {}
    dims = [in_dim, *hidden, out_dim]
    lead = (ens,) if ens else ()
    for i in range(len(dims) - 1):
        p[f"Dense_{i}"] = {
            "kernel": _glorot_uniform(gen, dims[i], dims[i + 1], (*lead, dims[i], dims[i + 1]), dtype),
            "bias": torch.zeros((*lead, dims[i + 1]), dtype=dtype),
        }
        if layer_norm and i < len(dims) - 2:
            p[f"LayerNorm_{i}"] = {
                "scale": torch.ones((*lead, dims[i + 1]), dtype=dtype),
                "bias": torch.zeros((*lead, dims[i + 1]), dtype=dtype),
            }
    return p


def init_params(seed: int, cfg: FQLConfig, dtype=torch.float32):
    """Deterministic synthetic initialisation with the upstream initialiser family (Glorot-uniform kernels,
    zero biases, LayerNorm scale 1 / bias 0, target_critic = critic). The random STREAM is torch's, not
    JAX's -- parity tests import these exact tensors into the CUDA population."""
    gen = torch.Generator().manual_seed(int(seed))
    O, A = cfg.ob_dim, cfg.action_dim
    critic = _init_mlp(gen, O + A, cfg.value_hidden_dims, 1, cfg.layer_norm, cfg.num_ensembles, dtype)
    params = {
        "critic": critic,
        "target_critic": clone_tree(critic),
        "actor_bc_flow": _init_mlp(gen, O + A + 1, cfg.actor_hidden_dims, A, cfg.actor_layer_norm, 0, dtype),
        "actor_onestep_flow": _init_mlp(gen, O + A, cfg.actor_hidden_dims, A, cfg.actor_layer_norm, 0, dtype),
    }
    return params


def clone_tree(t):
    if isinstance(t, dict):
        return {k: clone_tree(v) for k, v in t.items()}
    return t.clone()


def tree_map(f, *ts):
    if isinstance(ts[0], dict):
        return {k: tree_map(f, *[t[k] for t in ts]) for k in ts[0]}
    return f(*ts)


def tree_leaves(t, prefix=""):
    if isinstance(t, dict):
        out = []
        for k in t:
            out += tree_leaves(t[k], f"{prefix}/{k}" if prefix else k)
        return out
    return [(prefix, t)]


def tree_cast(t, dtype):
    return tree_map(lambda x: x.to(dtype), t)


def init_state(seed: int, cfg: FQLConfig, dtype=torch.float32):
    params = init_params(seed, cfg, dtype)
    zeros = lambda x: torch.zeros_like(x)
    trainable = {k: params[k] for k in ("critic", "actor_bc_flow", "actor_onestep_flow")}
    return {"params": params, "mu": tree_map(zeros, trainable), "nu": tree_map(zeros, trainable), "step": 0}


# --------------------------------------------------------------------------------------------------
# networks
# --------------------------------------------------------------------------------------------------
def _bf16(x):
    return x.to(torch.bfloat16).to(x.dtype)


class _QLinear(torch.autograd.Function):
    """y = bf16(x) @ bf16(W) with bf16-rounded operands in forward AND backward (fp accumulate)."""

    @staticmethod
    def forward(ctx, x, w):
        xq, wq = _bf16(x), _bf16(w)
        ctx.save_for_backward(xq, wq)
        return xq @ wq

    @staticmethod
    def backward(ctx, dy):
        xq, wq = ctx.saved_tensors
        dyq = _bf16(dy)
        return dyq @ wq.transpose(-1, -2), xq.transpose(-1, -2) @ dyq


class _QBias(torch.autograd.Function):
    """z = y + b; db is the column sum of the bf16-rounded dz (what the CUDA weight-gradient path sums)."""

    @staticmethod
    def forward(ctx, y, b):
        return y + b

    @staticmethod
    def backward(ctx, dz):
        return dz, _bf16(dz).sum(dim=tuple(range(dz.dim() - 1)))


class _QHead(torch.autograd.Function):
    """Head layer as the CUDA path computes it: forward on CUDA cores with the bf16 activations as stored
    and fp32 weights; backward through the tensor-core GEMMs, i.e. bf16(dy) against bf16(W) and bf16(x)."""

    @staticmethod
    def forward(ctx, x, w, b):
        xq = _bf16(x)
        ctx.save_for_backward(xq, w)
        return xq @ w + b

    @staticmethod
    def backward(ctx, dy):
        xq, w = ctx.saved_tensors
        dyq = _bf16(dy)
        return dyq @ _bf16(w).transpose(-1, -2), xq.transpose(-1, -2) @ dyq, dyq.sum(dim=tuple(range(dy.dim() - 1)))


def gelu_tanh(z):
    # flax nn.gelu default approximate=True
    return 0.5 * z * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (z + 0.044715 * z ** 3)))


def layer_norm(x, scale, bias, eps, park_bf16=False):
    # flax LayerNorm(use_fast_variance=True): var = E[x^2] - E[x]^2
    # park_bf16 (bf16 emulation of the CUDA chain kernels): the row statistics come from the fp32 activations, the
    # values that get normalised were parked as bf16 while the tensor core finished the layer.
    mean = x.mean(-1, keepdim=True)
    var = (x * x).mean(-1, keepdim=True) - mean * mean
    var = torch.clamp(var, min=0.0)
    if park_bf16:
        x = _bf16_ste(x)
    return (x - mean) * torch.rsqrt(var + eps) * scale + bias


def mlp_apply(p, x, n_layers, ln, eps, emulate_bf16=False, head_bf16=False):
    """MLP: [Dense -> GELU -> (LayerNorm)] x hidden, then Dense. x: [B, in]; params without ensemble axis.
    head_bf16: the fused rollout kernel runs the head on the tensor cores too (bf16 weights); the training
    step runs it on CUDA cores with fp32 weights."""
    for i in range(n_layers):
        W, b = p[f"Dense_{i}"]["kernel"], p[f"Dense_{i}"]["bias"]
        last = i == n_layers - 1
        if emulate_bf16 and (not last or head_bf16):
            x = _QBias.apply(_QLinear.apply(x, W), b)
        elif emulate_bf16 and last:
            x = _QHead.apply(x, W, b)
        else:
            x = x @ W + b
        if not last:
            x = gelu_tanh(x)
            if ln:
                x = layer_norm(x, p[f"LayerNorm_{i}"]["scale"], p[f"LayerNorm_{i}"]["bias"], eps,
                               park_bf16=emulate_bf16 and not head_bf16)  # the rollout kernel normalises fp32
    return x


class _Bf16STE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return _bf16(x)

    @staticmethod
    def backward(ctx, g):
        return g


def _bf16_ste(x):
    return _Bf16STE.apply(x)


def _ens_slice(p, e):
    return tree_map(lambda t: t[e], p)


def critic_apply(p, obs, act, cfg: FQLConfig, emulate_bf16=False):
    """Value net with E ensembles -> [E, B]."""
    x = torch.cat([obs, act], -1)
    n = len(cfg.value_hidden_dims) + 1
    outs = [mlp_apply(_ens_slice(p, e), x, n, cfg.layer_norm, cfg.ln_eps, emulate_bf16).squeeze(-1)
            for e in range(cfg.num_ensembles)]
    return torch.stack(outs, 0)


def actor_apply(p, obs, act, t, cfg: FQLConfig, emulate_bf16=False, head_bf16=False):
    xs = [obs, act] if t is None else [obs, act, t]
    n = len(cfg.actor_hidden_dims) + 1
    return mlp_apply(p, torch.cat(xs, -1), n, cfg.actor_layer_norm, cfg.ln_eps, emulate_bf16, head_bf16)


def sample_actions(params, obs, noise, cfg: FQLConfig, emulate_bf16=False, head_bf16=False):
    """FQLAgent.sample_actions: clip(pi_omega(obs, z), -1, 1); temperature is accepted upstream and unused."""
    with torch.no_grad():
        return torch.clamp(actor_apply(params["actor_onestep_flow"], obs, noise, None, cfg, emulate_bf16, head_bf16),
                           -1, 1)


def compute_flow_actions(params, obs, noise, cfg: FQLConfig, emulate_bf16=False):
    with torch.no_grad():
        a = noise
        F = cfg.flow_steps
        for i in range(F):
            t = torch.full((*obs.shape[:-1], 1), i / F, dtype=obs.dtype)
            a = a + actor_apply(params["actor_bc_flow"], obs, a, t, cfg, emulate_bf16) / F
        return torch.clamp(a, -1, 1)


# --------------------------------------------------------------------------------------------------
# losses / update
# --------------------------------------------------------------------------------------------------
def total_loss(params, batch, noise, cfg: FQLConfig, alpha: Optional[float] = None, emulate_bf16=False):
    """critic_loss + actor_loss and the info dict (keys as upstream: 'critic/...', 'actor/...').

    batch: observations [B,O], actions [B,A], next_observations [B,O], rewards [B], masks [B].
    noise: z_critic [B,A], x0 [B,A], t [B,1], z_actor [B,A], z_mse [B,A].
    Gradients flow into params['critic'], params['actor_bc_flow'], params['actor_onestep_flow'].
    """
    alpha = cfg.alpha if alpha is None else alpha
    obs, act, nobs = batch["observations"], batch["actions"], batch["next_observations"]
    rew, masks = batch["rewards"], batch["masks"]
    info = {}
    # ---- critic loss
    next_a = sample_actions(params, nobs, noise["z_critic"], cfg, emulate_bf16)
    with torch.no_grad():
        next_qs = critic_apply(params["target_critic"], nobs, next_a, cfg, emulate_bf16)
        next_q = next_qs.min(0).values if cfg.q_agg == "min" else next_qs.mean(0)
        target_q = rew + cfg.discount * masks * next_q
    q = critic_apply(params["critic"], obs, act, cfg, emulate_bf16)
    critic_loss = ((q - target_q) ** 2).mean()
    info.update({"critic/critic_loss": critic_loss, "critic/q_mean": q.mean(), "critic/q_max": q.max(),
                 "critic/q_min": q.min()})
    # ---- actor loss
    x0, t = noise["x0"], noise["t"]
    x_t = (1 - t) * x0 + t * act
    vel = act - x0
    pred = actor_apply(params["actor_bc_flow"], obs, x_t, t, cfg, emulate_bf16)
    bc_flow_loss = ((pred - vel) ** 2).mean()
    target_flow_actions = compute_flow_actions(params, obs, noise["z_actor"], cfg, emulate_bf16)
    actor_actions = actor_apply(params["actor_onestep_flow"], obs, noise["z_actor"], None, cfg, emulate_bf16)
    distill_loss = ((actor_actions - target_flow_actions) ** 2).mean()
    clipped = torch.clamp(actor_actions, -1, 1)
    frozen_critic = tree_map(lambda x: x.detach(), params["critic"])
    qs = critic_apply(frozen_critic, obs, clipped, cfg, emulate_bf16)
    qa = qs.mean(0)
    q_loss = -qa.mean()
    if cfg.normalize_q_loss:
        q_loss = q_loss * (1.0 / qa.abs().mean()).detach()
    actor_loss = bc_flow_loss + alpha * distill_loss + q_loss
    mse_actions = sample_actions(params, obs, noise["z_mse"], cfg, emulate_bf16)
    mse = ((mse_actions - act) ** 2).mean()
    info.update({"actor/actor_loss": actor_loss, "actor/bc_flow_loss": bc_flow_loss,
                 "actor/distill_loss": distill_loss, "actor/q_loss": q_loss, "actor/q": qa.mean(),
                 "actor/mse": mse})
    return critic_loss + actor_loss, info


def update(state, batch, noise, cfg: FQLConfig, alpha: Optional[float] = None, emulate_bf16=False,
           return_grads=False):
    """One FQLAgent.update: grads of total_loss, optax.adam(lr), Polyak target update. Returns (state, info)."""
    params = state["params"]
    trainable_names = ("critic", "actor_bc_flow", "actor_onestep_flow")
    work = {k: tree_map(lambda x: x.detach().clone().requires_grad_(k in trainable_names), params[k])
            for k in params}
    loss, info = total_loss(work, batch, noise, cfg, alpha, emulate_bf16)
    leaves = []
    for k in trainable_names:
        leaves += [t for _, t in tree_leaves(work[k])]
    grads_flat = torch.autograd.grad(loss, leaves, allow_unused=True)
    grads_flat = [torch.zeros_like(l) if g is None else g for g, l in zip(grads_flat, leaves)]
    it = iter(grads_flat)
    grads = {k: tree_map(lambda x: next(it), work[k]) for k in trainable_names}
    # grad statistics as upstream TrainState.apply_loss_fn: max/min over all leaves (target leaves are 0),
    # norm = sum over leaves of each leaf's L2 norm
    gmax = max([g.max() for g in grads_flat] + [torch.zeros(())])
    gmin = min([g.min() for g in grads_flat] + [torch.zeros(())])
    gnorm = sum(torch.linalg.vector_norm(g) for g in grads_flat)
    info = {k: float(v.detach()) for k, v in info.items()}
    info.update({"grad/max": float(gmax), "grad/min": float(gmin), "grad/norm": float(gnorm)})
    # optax.adam
    step = state["step"] + 1
    b1, b2, eps, lr = cfg.adam_b1, cfg.adam_b2, cfg.adam_eps, cfg.lr
    mu = tree_map(lambda m, g: b1 * m + (1 - b1) * g, state["mu"], grads)
    nu = tree_map(lambda v, g: b2 * v + (1 - b2) * g * g, state["nu"], grads)
    bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
    new_params = {}
    for k in trainable_names:
        new_params[k] = tree_map(lambda p, m, v: p.detach() - lr * (m / bc1) / (torch.sqrt(v / bc2) + eps),
                                 params[k], mu[k], nu[k])
    new_params["target_critic"] = tree_map(lambda p, tp: cfg.tau * p + (1 - cfg.tau) * tp,
                                           new_params["critic"], params["target_critic"])
    new_state = {"params": new_params, "mu": mu, "nu": nu, "step": step}
    if return_grads:
        return new_state, info, grads
    return new_state, info


# --------------------------------------------------------------------------------------------------
# world model + simulated evaluation
# --------------------------------------------------------------------------------------------------
def init_envmodel(seed: int, ob_dim: int, action_dim: int, hidden=(128, 256, 128), dtype=torch.float32):
    """flax defaults for nn.Dense(h) without initialiser (envmodel/baseline.py:34-35): LeCun-normal kernels,
    zero biases; LayerNorm scale 1 / bias 0. Param names follow flax auto-naming."""
    gen = torch.Generator().manual_seed(int(seed))

    def lecun(fi, fo):
        return (torch.randn((fi, fo), generator=gen, dtype=torch.float64) / math.sqrt(fi)).to(dtype)

    def mlp(in_dim, out_dim):
        dims = [in_dim, *hidden, out_dim]
        return {f"Dense_{i}": {"kernel": lecun(dims[i], dims[i + 1]), "bias": torch.zeros(dims[i + 1], dtype=dtype)}
                for i in range(len(dims) - 1)}

    state = mlp(ob_dim + action_dim, ob_dim)
    state["LayerNorm_0"] = {"scale": torch.ones(ob_dim + action_dim, dtype=dtype),
                            "bias": torch.zeros(ob_dim + action_dim, dtype=dtype)}
    term = mlp(ob_dim, 1)
    return {"state": state, "termination": term}


def _dense(x, d, emulate_bf16):
    if emulate_bf16:  # bf16 operands, wide accumulation, fp32 bias -- as the tensor-core path
        return _bf16(x) @ _bf16(d["kernel"]) + d["bias"]
    return x @ d["kernel"] + d["bias"]


def state_predictor_apply(p, obs, act, eps=1e-6, emulate_bf16=False):
    """BaselineStatePredictor.__call__ (/root/reference/envmodel/baseline.py:25-37)."""
    x = torch.cat([obs, act], -1)
    x = layer_norm(x, p["LayerNorm_0"]["scale"], p["LayerNorm_0"]["bias"], eps)
    n = sum(1 for k in p if k.startswith("Dense_"))
    for i in range(n - 1):
        x = torch.relu(_dense(x, p[f"Dense_{i}"], emulate_bf16))
    return _dense(x, p[f"Dense_{n - 1}"], emulate_bf16) + obs


def termination_predictor_apply(p, obs, emulate_bf16=False):
    """TerminationPredictor.__call__ with train=False (dropout is the identity)
    (/root/reference/envmodel/termination_predictor.py:14-21)."""
    x = obs
    n = sum(1 for k in p if k.startswith("Dense_"))
    for i in range(n - 1):
        x = torch.relu(_dense(x, p[f"Dense_{i}"], emulate_bf16))
    return _dense(x, p[f"Dense_{n - 1}"], emulate_bf16).squeeze(-1)


def rollout(agent_params, envmodel, init_obs, policy_noise, cfg: FQLConfig, max_episode_steps: int,
            emulate_bf16=False, return_trajectory=False, horizon: int = 0):
    """Simulated evaluation: evaluate_actor_fn (/root/reference/evaluator/evaluation.py:67-114) driving
    OfflineTaskWithSimulatedEvaluations.step (/root/reference/task/offline_task_simulated.py:85-107).

    init_obs [n,O]; policy_noise [T,n,A] (one draw per lockstep step); horizon > 0 stops after that many
    lockstep steps (fixed-horizon parity tests; unfinished episodes keep length 0). Returns dict with per-episode
    success (bool), length (int, the step of the first termination/truncation), the scalar success rate
    (= stats['success'] of the reference) and the number of lockstep steps executed.
    """
    n = init_obs.shape[0]
    obs = init_obs.clone()
    done_once = torch.zeros(n, dtype=torch.bool)  # == invalidate[i] of the reference
    success = torch.zeros(n, dtype=torch.bool)
    length = torch.zeros(n, dtype=torch.int64)
    traj = []
    steps = 0
    with torch.no_grad():
        while True:
            a = sample_actions(agent_params, obs, policy_noise[steps], cfg, emulate_bf16, head_bf16=True)
            a = torch.clamp(a, -1, 1)  # evaluation.py:96
            nxt = state_predictor_apply(envmodel["state"], obs, a, emulate_bf16=emulate_bf16)
            logit = termination_predictor_apply(envmodel["termination"], nxt, emulate_bf16=emulate_bf16)
            term = logit > 0
            steps += 1
            trunc = torch.full_like(term, steps >= max_episode_steps)
            first = (term | trunc) & ~done_once
            success |= first & term
            length = torch.where(first, torch.full_like(length, steps), length)
            if return_trajectory:
                traj.append((obs.clone(), a.clone(), done_once.clone(), logit.clone()))
            done_now = term | trunc | done_once  # evaluation.py:100
            done_once |= term | trunc            # offline_task_simulated.py:102-105
            obs = nxt                            # offline_task_simulated.py:93 (all envs, finished or not)
            if bool(done_now.all()) or (horizon and steps >= horizon):
                break
    run_len = torch.where(length > 0, length, torch.full_like(length, steps))
    out = {"success": success, "length": length, "success_rate": float(success.float().mean()),
           "steps": steps, "returns": -run_len.to(torch.float32) + success.to(torch.float32)}
    if return_trajectory:
        out["trajectory"] = traj
    return out
