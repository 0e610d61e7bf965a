"""Generates tests/golden/successive_halving.json by EXECUTING the reference's own
/root/reference/hpo/successive_halving.py (with a stub of trainer.config, whose reference version does not import on
Python >= 3.11). Run in the build container only; the JSON is what the tests read.

    python tests/golden/make_successive_halving.py
"""
import csv
import importlib.util
import json
import os
import sys
import types
from dataclasses import dataclass

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


@dataclass(frozen=True)
class ExperimentConfig:
    seed: int | None = None
    alpha: float | None = None


def load_reference():
    tc = types.ModuleType("trainer.config")
    tc.ExperimentConfig = ExperimentConfig
    tr = types.ModuleType("trainer")
    tr.config = tc
    sys.modules["trainer"], sys.modules["trainer.config"] = tr, tc
    hp = types.ModuleType("hpo")
    hp.__path__ = [os.path.join(REF, "hpo")]
    sys.modules["hpo"] = hp
    for name in ("strategy", "successive_halving"):
        spec = importlib.util.spec_from_file_location(f"hpo.{name}", os.path.join(REF, "hpo", f"{name}.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[f"hpo.{name}"] = mod
        spec.loader.exec_module(mod)
    return sys.modules["hpo.successive_halving"].SuccessiveHalving


def logspace(n):
    import numpy as np
    return np.logspace(np.log10(3), np.log10(1000), num=n).tolist()


def run_case(SH, population, total, f, h, score_fn, n_evals):
    import math
    sh = SH(population=set(population), total_evaluations=total, fraction=f, history_length=h)
    trace = []
    current = sh.sample()
    for ev in range(1, n_evals + 1):
        for c in sorted(current, key=lambda c: (c.alpha, c.seed)):
            sh.update(c, score_fn(c, ev))
        current = sh.sample()
        trace.append(sorted([[c.seed, c.alpha] for c in current]))
    return {"milestones": sh.halving_milestones, "trace": trace}


def main():
    import math
    SH = load_reference()
    cases = []
    analytic = lambda c, ev: -abs(math.log10(c.alpha) - 2.0)
    for (n, total, f, h) in [(64, 50, 0.5, 2), (40, 50, 0.5, 2), (40, 50, 0.25, 1), (40, 50, 0.75, 8), (16, 10, 0.5, 1),
                             (256, 50, 0.5, 2), (64, 50, 0.5, 1), (20, 50, 0.5, 3)]:
        pop = [ExperimentConfig(seed=0, alpha=a) for a in logspace(n)]
        out = run_case(SH, pop, total, f, h, analytic, total)
        cases.append({"kind": "analytic", "n": n, "total": total, "fraction": f, "history": h, **out})
    # a score that decays for the leaders, so eliminated candidates re-enter (the reference ranks frozen histories too)
    decay = lambda c, ev: (1.0 / ev if c.alpha > 100 else 0.3) - 1e-3 * math.log10(c.alpha)
    pop = [ExperimentConfig(seed=0, alpha=a) for a in logspace(16)]
    cases.append({"kind": "decay", "n": 16, "total": 20, "fraction": 0.5, "history": 1,
                  **run_case(SH, pop, 20, 0.5, 1, decay, 20)})
    # replay of the reference's logged real success curves (results/real_success_rates_*.csv: seed,alpha,step,success)
    for env in ("cube", "antsoccer"):
        rows = list(csv.DictReader(open(os.path.join(REF, "results", f"real_success_rates_{env}.csv"))))
        curves = {}
        for r in rows:
            curves.setdefault((int(r["seed"]), float(r["alpha"])), []).append((int(r["step"]), float(r["success"])))
        curves = {k: [s for _, s in sorted(v)] for k, v in curves.items()}
        keys = sorted(curves)
        pop = [ExperimentConfig(seed=s, alpha=a) for s, a in keys]
        n_evals = min(len(v) for v in curves.values())
        score = lambda c, ev: curves[(c.seed, c.alpha)][ev - 1]
        for (f, h) in [(0.5, 2), (0.25, 1)]:
            out = run_case(SH, pop, n_evals, f, h, score, n_evals)
            cases.append({"kind": f"csv:{env}", "n": len(pop), "total": n_evals, "fraction": f, "history": h,
                          "population": [[s, a] for s, a in keys], "curves": [curves[k][:n_evals] for k in keys], **out})
    with open(os.path.join(HERE, "successive_halving.json"), "w") as fh:
        json.dump({"source": "executed /root/reference/hpo/successive_halving.py", "cases": cases}, fh)
    print("wrote", len(cases), "cases")


if __name__ == "__main__":
    main()
