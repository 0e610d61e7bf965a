"""Loss functions of world-model training with the reference's names and signatures (/root/reference/envmodel/loss.py:7-111).
The trainers do not call these on the training path: they read WHICH loss (and which weights, from the functools.partial
the reference builds at train_env_model.py:40-45,63-68,81) was configured and select the matching device kernel
(wm_state_loss_kernel / wm_focal_loss_kernel). The numpy bodies below serve host-side analysis of logged predictions."""
from typing import Dict, Tuple

import numpy as np


def _bce(logits, labels):
    return np.maximum(logits, 0) - logits * labels + np.log1p(np.exp(-np.abs(logits)))


def mean_squared_error(predictions, targets):
    return np.mean(np.square(np.asarray(predictions) - np.asarray(targets)))


def weighted_binary_cross_entropy(predictions, targets, true_weight: float = 1.0) -> Tuple[float, Dict[str, float]]:
    predictions, targets = np.asarray(predictions, np.float32), np.asarray(targets)
    loss = _bce(predictions, targets.astype(np.float32))
    true_loss, false_loss = np.where(targets == 1, loss, 0.0), np.where(targets == 0, loss, 0.0)
    total = np.mean((true_weight * true_loss + false_loss) / (true_weight + 1))
    with np.errstate(invalid="ignore", divide="ignore"):
        logs = {"true_loss": np.sum(true_loss) / np.float32(np.sum(targets == 1)),
                "false_loss": np.sum(false_loss) / np.float32(np.sum(targets == 0))}
    return total, logs


def focal_loss(predictions, targets, alpha: float = 0.25, gamma: float = 2.0) -> Tuple[float, Dict[str, float]]:
    predictions, t = np.asarray(predictions, np.float32), np.asarray(targets).astype(np.float32)
    p = 1.0 / (1.0 + np.exp(-predictions))
    ce = _bce(predictions, t)
    pt = np.where(t == 1, p, 1 - p)
    loss = np.mean(np.where(t == 1, alpha, 1 - alpha) * (1.0 - pt) ** gamma * ce)
    nt, nf = np.float32(np.sum(t == 1)), np.float32(np.sum(t == 0))
    return loss, {"true_loss": loss * nt / (nt + np.float32(1e-8)), "false_loss": loss * nf / (nf + np.float32(1e-8))}


def state_prediction_loss(predicted_next_observations, next_observations, predicted_termination, terminations,
                          reconstructed_observations, observations, true_termination_weight: float,
                          termination_weight: float, reconstruction_weight: float):
    nol = mean_squared_error(predicted_next_observations, next_observations)
    tl, tlogs = (weighted_binary_cross_entropy(predicted_termination, terminations, true_termination_weight)
                 if termination_weight > 0 else (0.0, {}))
    rl = mean_squared_error(reconstructed_observations, observations) if reconstruction_weight > 0 else 0.0
    loss = (nol + termination_weight * tl + reconstruction_weight * rl) / (1 + termination_weight + reconstruction_weight)
    logs = {"next_observation_loss": nol, "loss": loss}
    if termination_weight > 0:
        logs.update(termination_loss=tl, true_termination_loss=tlogs["true_loss"], false_termination_loss=tlogs["false_loss"])
    if reconstruction_weight > 0:
        logs["reconstruction_loss"] = rl
    return loss, logs
