"""CPU oracle for world-model training.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module, and only as the checker /
the timed CPU baseline; the product path never imports it.

What it restates (every function cites the reference lines it follows):
  * losses                     /root/reference/envmodel/loss.py:7-111
  * state-predictor step       /root/reference/envmodel/state_predictor_trainer.py:70-125
  * termination-predictor step /root/reference/envmodel/termination_predictor_trainer.py:58-126
  * initial-observation VAE    /root/reference/envmodel/initial_observation.py:9-78 (model + vae_loss; no trainer there)
  * optimiser                  optax.adam(optax.cosine_decay_schedule(init_learning_rate, steps))
                               (state_predictor_trainer.py:57-61, termination_predictor_trainer.py:46-50);
                               optax is a third-party dependency absent here -- its published algorithm: Adam with
                               b1=0.9, b2=0.999, eps=1e-8 outside the root, bias correction with count+1, learning
                               rate lr0 * 0.5 * (1 + cos(pi * min(count, T) / T)) read before the count increments.

PARITY UNPINNED: the reference holds no tests or golden vectors for this path and jax/flax/optax cannot be imported
here. The backward pass is torch autograd, i.e. an independent derivation of what the CUDA kernels compute by hand.
The dropout mask of the termination predictor is an explicit argument (JAX threefry bits cannot be reproduced); the
reference passes the same rng every step (termination_predictor_trainer.py:69), i.e. one fixed mask.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

from .fql_oracle import state_predictor_apply, termination_predictor_apply


def mean_squared_error(pred, target):
    """envmodel/loss.py:7-9"""
    return torch.mean((pred - target) ** 2)


def sigmoid_bce(logits, labels):
    """optax.sigmoid_binary_cross_entropy: -labels*log_sigmoid(x) - (1-labels)*log_sigmoid(-x)."""
    return -labels * F.logsigmoid(logits) - (1.0 - labels) * F.logsigmoid(-logits)


def weighted_binary_cross_entropy(pred, targets, true_weight=1.0):
    """envmodel/loss.py:12-28 (targets: bool)"""
    t = targets.to(pred.dtype)
    loss = sigmoid_bce(pred, t)
    true_loss = torch.where(targets, loss, torch.zeros_like(loss))
    false_loss = torch.where(~targets, loss, torch.zeros_like(loss))
    total = torch.mean((true_weight * true_loss + false_loss) / (true_weight + 1))
    logs = {"true_loss": true_loss.sum() / targets.sum(), "false_loss": false_loss.sum() / (~targets).sum()}
    return total, logs


def focal_loss(pred, targets, alpha=0.25, gamma=2.0):
    """envmodel/loss.py:31-62 (the logs scale the MEAN loss by the class counts, as written there)"""
    t = targets.to(pred.dtype)
    p = torch.sigmoid(pred)
    ce = sigmoid_bce(pred, t)
    pt = torch.where(targets, p, 1 - p)
    loss = torch.mean(torch.where(targets, alpha, 1 - alpha) * (1.0 - pt) ** gamma * ce)
    nt, nf = targets.sum(), (~targets).sum()
    logs = {"true_loss": loss * nt / (nt + 1e-8), "false_loss": loss * nf / (nf + 1e-8)}
    return loss, logs


def state_prediction_loss(pred_next, next_obs, pred_term, terminations, recon, obs, true_termination_weight,
                          termination_weight, reconstruction_weight):
    """envmodel/loss.py:65-111"""
    nol = mean_squared_error(pred_next, next_obs)
    if termination_weight > 0:
        tl, tlogs = weighted_binary_cross_entropy(pred_term, terminations, true_termination_weight)
    else:
        tl, tlogs = 0.0, {}
    rl = mean_squared_error(recon, obs) if reconstruction_weight > 0 else 0.0
    loss = (nol + termination_weight * tl + reconstruction_weight * rl) / (1 + termination_weight + reconstruction_weight)
    logs = {"next_observation_loss": nol, "loss": loss}
    if termination_weight > 0:
        logs["termination_loss"] = tl
        logs["true_termination_loss"] = tlogs["true_loss"]
        logs["false_termination_loss"] = tlogs["false_loss"]
    if reconstruction_weight > 0:
        logs["reconstruction_loss"] = rl
    return loss, logs


def cosine_lr(count: int, lr0: float, decay_steps: int) -> float:
    """optax.cosine_decay_schedule(lr0, decay_steps) with alpha=0, exponent=1."""
    return lr0 * 0.5 * (1.0 + math.cos(math.pi * min(count, decay_steps) / decay_steps))


def tree_leaves(p, prefix=""):
    for k in sorted(p):
        if isinstance(p[k], dict):
            yield from tree_leaves(p[k], prefix + k + "/")
        else:
            yield prefix + k, p[k]


def new_opt_state(params):
    z = lambda t: {k: (z(v) if isinstance(v, dict) else torch.zeros_like(v)) for k, v in t.items()}
    return {"count": 0, "m": z(params), "v": z(params)}


def _adam(params, grads, opt, lr0, decay_steps):
    lr = cosine_lr(opt["count"], lr0, decay_steps)
    c = opt["count"] + 1
    bc1, bc2 = 1 - 0.9 ** c, 1 - 0.999 ** c

    def walk(p, g, m, v):
        for k in p:
            if isinstance(p[k], dict):
                walk(p[k], g[k], m[k], v[k])
            else:
                m[k] = 0.9 * m[k] + 0.1 * g[k]
                v[k] = 0.999 * v[k] + 0.001 * g[k] ** 2
                p[k] = p[k] - lr * (m[k] / bc1) / (torch.sqrt(v[k] / bc2) + 1e-8)

    walk(params, grads, opt["m"], opt["v"])
    opt["count"] = c
    return lr


def _with_grad(params):
    return {k: (_with_grad(v) if isinstance(v, dict) else v.clone().requires_grad_(True)) for k, v in params.items()}


def _grads(params):
    return {k: (_grads(v) if isinstance(v, dict) else v.grad) for k, v in params.items()}


def _detach(params):
    return {k: (_detach(v) if isinstance(v, dict) else v.detach()) for k, v in params.items()}


def state_step(params, opt, batch: Dict[str, torch.Tensor], term_params: Optional[dict], *, lr0, decay_steps,
               true_termination_weight=30.0, termination_weight=1.0, reconstruction_weight=0.0, train=True):
    """StatePredictorTrainer.train_step / eval_step (state_predictor_trainer.py:70-125): the termination predictor is
    frozen and applied to the PREDICTED next observation with train=False; targets are rewards == 0.
    Returns (new params, logs, grads)."""
    p = _with_grad(params) if train else params
    pred = state_predictor_apply(p, batch["observations"], batch["actions"])
    logit = termination_predictor_apply(term_params, pred) if termination_weight > 0 else None
    loss, logs = state_prediction_loss(pred, batch["next_observations"], logit, batch["rewards"] == 0,
                                       batch["observations"], batch["observations"], true_termination_weight,
                                       termination_weight, reconstruction_weight)
    logs = {k: float(v.detach()) if torch.is_tensor(v) else float(v) for k, v in logs.items()}
    if not train:
        return params, logs, None
    loss.backward()
    g = _grads(p)
    new = _detach(p)
    logs["learning_rate"] = _adam(new, g, opt, lr0, decay_steps)
    return new, logs, g


def termination_step(params, opt, batch, dropout_mask=None, *, lr0, decay_steps, alpha=0.25, gamma=2.0, train=True):
    """TerminationPredictorTrainer.train_step / eval_step (termination_predictor_trainer.py:58-126): input is
    next_observations, under dropout (explicit multiplier) when training; eval adds accuracy / precision / recall / f1."""
    p = _with_grad(params) if train else params
    x = batch["next_observations"]
    if train and dropout_mask is not None:
        x = x * dropout_mask
    logit = termination_predictor_apply(p, x)
    targets = batch["rewards"] == 0
    loss, logs = focal_loss(logit, targets, alpha, gamma)
    logs = {k: float(v.detach()) if torch.is_tensor(v) else float(v) for k, v in logs.items()}
    logs["loss"] = float(loss.detach())
    if not train:
        pred = logit > 0
        tp, pp, ap = float((pred & targets).sum()), float(pred.sum()), float(targets.sum())
        prec = tp / pp if pp > 0 else 1.0
        rec = tp / ap if ap > 0 else 1.0
        logs.update(accuracy=float((pred == targets).to(torch.float64).mean()), precision=prec, recall=rec,
                    f1=(2 * prec * rec / (prec + rec)) if (prec + rec) > 0 else 0.0)
        return params, logs, None
    loss.backward()
    g = _grads(p)
    new = _detach(p)
    logs["learning_rate"] = _adam(new, g, opt, lr0, decay_steps)
    return new, logs, g


def multistep_step(params, opt, batch, term_params, *, lr0, decay_steps, true_termination_weight=30.0,
                   termination_weight=1.0, reconstruction_weight=0.0, train=True):
    """StatePredictorTrainer.train_step with the MultistepStatePredictor (envmodel/multistep.py:34-54): the baseline cell
    scanned along axis 1 of actions from observations[:, 0, :], each prediction fed back as the next observation; the loss
    sees all T predictions [B, T, O] (and the termination predictor's logits on them) against next_observations / rewards
    [B, T]. `params` is the cell's tree (params["ScanCell_0"]["cell"] in the reference's checkpoint)."""
    p = _with_grad(params) if train else params
    o = batch["observations"][:, 0, :]
    preds, recon = [], []
    for t in range(batch["actions"].shape[1]):
        recon.append(o)
        o = state_predictor_apply(p, o, batch["actions"][:, t, :])
        preds.append(o)
    pred, recon = torch.stack(preds, 1), torch.stack(recon, 1)
    logit = termination_predictor_apply(term_params, pred) if termination_weight > 0 else None
    loss, logs = state_prediction_loss(pred, batch["next_observations"], logit, batch["rewards"] == 0, recon,
                                       batch["observations"], true_termination_weight, termination_weight,
                                       reconstruction_weight)
    logs = {k: float(v.detach()) if torch.is_tensor(v) else float(v) for k, v in logs.items()}
    if not train:
        return params, logs, None
    loss.backward()
    g = _grads(p)
    new = _detach(p)
    logs["learning_rate"] = _adam(new, g, opt, lr0, decay_steps)
    return new, logs, g


# ---- initial-observation VAE (/root/reference/envmodel/initial_observation.py:9-78) ----------------------------------
def vae_encoder_apply(p, x):
    """Encoder.__call__ (initial_observation.py:13-18): [Dense, relu] per hidden dim, then the mean head (the next
    Dense flax names) and the logvar head (the one after)."""
    n = sum(1 for k in p if k.startswith("Dense_"))
    for i in range(n - 2):
        x = torch.relu(x @ p[f"Dense_{i}"]["kernel"] + p[f"Dense_{i}"]["bias"])
    mean = x @ p[f"Dense_{n - 2}"]["kernel"] + p[f"Dense_{n - 2}"]["bias"]
    logvar = x @ p[f"Dense_{n - 1}"]["kernel"] + p[f"Dense_{n - 1}"]["bias"]
    return mean, logvar


def vae_decoder_apply(p, z):
    """Decoder.__call__ (initial_observation.py:25-29)."""
    n = sum(1 for k in p if k.startswith("Dense_"))
    for i in range(n - 1):
        z = torch.relu(z @ p[f"Dense_{i}"]["kernel"] + p[f"Dense_{i}"]["bias"])
    return z @ p[f"Dense_{n - 1}"]["kernel"] + p[f"Dense_{n - 1}"]["bias"]


def vae_loss(params, observations, eps, reconstruction_weight):
    """InitialObservationEnvModel.__call__ + vae_loss (initial_observation.py:41-78). `eps` is the reparameterisation
    noise jax.random.normal(key, std.shape) of :43, an explicit argument here (JAX threefry bits cannot be reproduced)."""
    mu, logvar = vae_encoder_apply(params["encoder"], observations)
    z = mu + eps * torch.exp(0.5 * logvar)
    recon = vae_decoder_apply(params["decoder"], z)
    reconstruction_loss = torch.mean((observations - recon) ** 2)
    kl = -0.5 * torch.mean(1 + logvar - mu ** 2 - torch.exp(logvar))
    loss = (reconstruction_weight * reconstruction_loss + kl) / (reconstruction_weight + 1)
    return loss, {"reconstruction_loss": reconstruction_loss, "kl": kl, "loss": loss}


def vae_step(params, opt, observations, eps, *, reconstruction_weight, lr0, decay_steps, train=True):
    """One optimiser step on vae_loss. The reference has no trainer for this model; the optimiser is the one of its other
    world-model trainers (state_predictor_trainer.py:57-61). Returns (new params, logs, grads)."""
    p = _with_grad(params) if train else params
    loss, logs = vae_loss(p, observations, eps, reconstruction_weight)
    logs = {k: float(v.detach()) for k, v in logs.items()}
    if not train:
        return params, logs, None
    loss.backward()
    g = _grads(p)
    new = _detach(p)
    logs["learning_rate"] = _adam(new, g, opt, lr0, decay_steps)
    return new, logs, g
