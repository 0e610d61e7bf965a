"""Generates tests/golden/strategy_evaluator.json by EXECUTING the reference's own
/root/reference/evaluator/{experiment_from_file,strategy_evaluator}.py + hpo/successive_halving.py over synthetic
eval.csv files (stubs: trainer.config -- the reference's does not import on Python >= 3.11 --, matplotlib, seaborn,
utils.tasks). Run in the build container only; the JSON (inputs and the reference's outputs) is what the tests read.

    python tests/golden/make_strategy_evaluator.py
"""
import importlib.util
import json
import os
import sys
import tempfile
import types
from dataclasses import dataclass
from pathlib import Path

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


@dataclass(frozen=True)
class ExperimentConfig:
    seed: int | None = None
    alpha: float | None = None


@dataclass
class TrainerConfig:
    steps: int = 1000000
    eval_interval: int = 100000
    eval_episodes: int = 50
    save_directory: Path = Path("exp/")
    env_name: str = "cube-single-play-singletask-task2-v0"


def _load(pkg, name):
    spec = importlib.util.spec_from_file_location(f"{pkg}.{name}", os.path.join(REF, pkg, f"{name}.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[f"{pkg}.{name}"] = mod
    spec.loader.exec_module(mod)
    return mod


def load_reference():
    tc = types.ModuleType("trainer.config")
    tc.ExperimentConfig, tc.TrainerConfig = ExperimentConfig, TrainerConfig
    tr = types.ModuleType("trainer"); tr.config = tc
    sys.modules["trainer"], sys.modules["trainer.config"] = tr, tc
    for stub in ("matplotlib", "matplotlib.pyplot", "seaborn"):
        sys.modules[stub] = types.ModuleType(stub)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    ut = types.ModuleType("utils"); ut.__path__ = []
    tk = types.ModuleType("utils.tasks")
    tk.get_task_filename = lambda e: e
    tk.get_task_title = lambda e: e
    sys.modules["utils"], sys.modules["utils.tasks"] = ut, tk
    for pkg in ("hpo", "evaluator"):
        m = types.ModuleType(pkg); m.__path__ = [os.path.join(REF, pkg)]
        sys.modules[pkg] = m
    _load("hpo", "strategy")
    sh = _load("hpo", "successive_halving").SuccessiveHalving
    eff = _load("evaluator", "experiment_from_file")
    se = _load("evaluator", "strategy_evaluator").StrategyEvaluator
    return sh, eff, se


def curves(rng, configs, n_evals, kind):
    out = {}
    for i, c in enumerate(configs):
        if kind == "monotone":       # saturating curves with different plateaus
            plateau = rng.uniform(0.1, 1.0)
            y = plateau * (1 - np.exp(-np.arange(1, n_evals + 1) / rng.uniform(1, 4)))
        elif kind == "noisy":        # success rates of 50 episodes: multiples of 0.02 with noise, ties likely
            base = rng.uniform(0.0, 0.9)
            y = np.clip(base + 0.15 * rng.standard_normal(n_evals), 0, 1)
            y = np.round(y * 50) / 50
        else:                        # all zeros except a late riser
            y = np.zeros(n_evals)
            if i == len(configs) - 1:
                y[n_evals // 2:] = 0.4
        out[c] = [float(v) for v in y]
    return out


def main():
    SH, eff, SE = load_reference()
    rng = np.random.default_rng(0)
    cases = []
    specs = [(8, 10, 0.5, 1, "monotone"), (16, 10, 0.5, 2, "noisy"), (12, 20, 0.25, 3, "noisy"), (6, 10, 0.5, 1, "flat"),
             (32, 10, 0.75, 1, "monotone")]
    for n, n_evals, f, h, kind in specs:
        alphas = np.logspace(np.log10(3), np.log10(1000), num=n).tolist()
        configs = [ExperimentConfig(seed=int(rng.integers(0, 10000)), alpha=a) for a in alphas]
        cv = curves(rng, configs, n_evals, kind)
        cfg = TrainerConfig(steps=n_evals * 1000, eval_interval=1000)
        with tempfile.TemporaryDirectory() as tmp:
            cfg.save_directory = Path(tmp)
            for j, c in enumerate(configs):
                # two run directories per config: the newer timestamp must win
                for stamp, scale in (("20240101_000000", 0.0), ("20250607_101112", 1.0)):
                    d = Path(tmp) / cfg.env_name / f"{eff.config_to_name(c)}_{stamp}"
                    d.mkdir(parents=True)
                    with open(d / "eval.csv", "w") as fh:
                        fh.write("success,step\n")
                        for k, v in enumerate(cv[c]):
                            fh.write(f"{v * scale},{(k + 1) * 1000}\n")
            total = n_evals * n
            strat = SH(population=set(configs), total_evaluations=total, fraction=f, history_length=h)
            ev = SE(strat, cfg)
            curve = ev.evaluate()
            stopped = {eff.config_to_name(c): e.current_step for c, e in ev.experiments.items()}
        cases.append({"n": n, "n_evals": n_evals, "fraction": f, "history_length": h, "kind": kind,
                      "total_evaluations": total, "eval_interval": 1000, "steps": n_evals * 1000,
                      "configs": [[c.seed, c.alpha] for c in configs], "curves": [cv[c] for c in configs],
                      "score_curve": {str(k): v for k, v in curve.items()}, "stopped_at": stopped,
                      "names": [eff.config_to_name(c) for c in configs]})
    with open(os.path.join(HERE, "strategy_evaluator.json"), "w") as f:
        json.dump({"generated_by": "tests/golden/make_strategy_evaluator.py executing /root/reference/evaluator + hpo", "cases": cases}, f)
    print("wrote", len(cases), "cases;", [len(c["score_curve"]) for c in cases])


if __name__ == "__main__":
    main()
