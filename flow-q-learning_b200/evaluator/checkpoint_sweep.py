"""Simulated success rates of MANY saved agents at once, and their agreement with logged real success rates -- the
measurement loop of the reference's /root/reference/evaluate_success_rate_correlation.py:43-66,84-93, where every row
(seed, alpha, step) loads `seed_<seed>_alpha_<alpha>_*/checkpoint_<step>.pkl` and runs one simulated evaluation.
Here up to `batch` checkpoints are loaded into the slots of ONE device population and scored by ONE fused rollout launch
(fqlpop_rollout), instead of one evaluation per checkpoint. The Gaussian-process acquisition loop and the plots of that
script are analysis code outside the hot path and are not reproduced."""
from pathlib import Path
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import yaml

from utils.agent import load_agent


def find_experiment(checkpoint_root: Path, env_name: str, seed: int, alpha: str) -> Optional[Path]:
    """First directory matching the reference's naming, or None (the reference prints a message and skips the row)."""
    return next(iter(sorted((Path(checkpoint_root) / env_name).glob(f"seed_{seed}_alpha_{alpha}_*"))), None)


def evaluate_checkpoints(task, checkpoints: Sequence[Tuple[Path, str]], batch: int = 16, seed: Optional[int] = None,
                         device=None) -> List[float]:
    """checkpoints: (agent_directory, agent_filename without extension). Returns stats["success"] of the simulated
    evaluation of each, in order. All agents of a call must share their network shapes (config.yaml)."""
    from fqlpop.population import Population  # needs libfqlpop.so; no CPU fallback
    checkpoints = [(Path(d), f) for d, f in checkpoints]
    if not checkpoints:
        return []
    example = task.sample("train", 1)
    hyper = yaml.unsafe_load((checkpoints[0][0] / "config.yaml").read_text())
    hyper = dict(hyper)
    hyper["ob_dims"] = tuple(np.shape(example["observations"])[-1:])
    hyper["action_dim"] = int(np.shape(example["actions"])[-1])
    slots = min(batch, len(checkpoints))
    population = Population(hyper, [float(hyper.get("alpha", 10.0))] * slots, list(range(slots)),
                            device=device or getattr(task, "device", None))
    scores: List[float] = []
    try:
        for first in range(0, len(checkpoints), slots):
            chunk = checkpoints[first:first + slots]
            for slot, (directory, filename) in enumerate(chunk):
                load_agent(directory, example, agent_filename=filename, population=population, slot=slot)
            population.set_active(list(range(len(chunk))))
            out = task.rollout_population(population, seed=seed)
            rates = out["success_rate"].cpu().numpy()
            scores.extend(float(rates[s]) for s in range(len(chunk)))
    finally:
        population.close()
    return scores


def success_rate_correlation(real: Sequence[float], simulated: Sequence[float]) -> Dict[str, float]:
    """Pearson r, Spearman rho and Kendall tau with their p-values, as reported in the reference's figure title
    (evaluate_success_rate_correlation.py:88-90)."""
    from scipy.stats import kendalltau, pearsonr, spearmanr
    x, y = np.asarray(real, dtype=np.float64), np.asarray(simulated, dtype=np.float64)
    r, p = pearsonr(x, y)
    rho, rho_p = spearmanr(x, y)
    tau, tau_p = kendalltau(x, y)
    return {"pearson_r": float(r), "pearson_p": float(p), "spearman_rho": float(rho), "spearman_p": float(rho_p),
            "kendall_tau": float(tau), "kendall_p": float(tau_p)}
