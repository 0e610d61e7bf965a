"""SuccessiveHalving / Identity against golden vectors produced by EXECUTING the reference's
/root/reference/hpo/successive_halving.py (tests/golden/make_successive_halving.py)."""
import json
import math
import os

import pytest

from hpo.identity import Identity
from hpo.successive_halving import SuccessiveHalving
from trainer.config import ExperimentConfig

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "successive_halving.json")))


def _score_fn(case):
    if case["kind"] == "analytic":
        return lambda c, ev: -abs(math.log10(c.alpha) - 2.0)
    if case["kind"] == "decay":
        return lambda c, ev: (1.0 / ev if c.alpha > 100 else 0.3) - 1e-3 * math.log10(c.alpha)
    curves = {(s, a): cur for (s, a), cur in zip(map(tuple, case["population"]), case["curves"])}
    return lambda c, ev: curves[(c.seed, c.alpha)][ev - 1]


def _population(case):
    if "population" in case:
        return [ExperimentConfig(seed=int(s), alpha=a) for s, a in case["population"]]
    import numpy as np
    return [ExperimentConfig(seed=0, alpha=a) for a in np.logspace(np.log10(3), np.log10(1000), num=case["n"]).tolist()]


@pytest.mark.parametrize("case", GOLD["cases"], ids=lambda c: f"{c['kind']}-n{c['n']}-f{c['fraction']}-h{c['history']}")
def test_successive_halving_matches_reference_trace(case):
    sh = SuccessiveHalving(population=set(_population(case)), total_evaluations=case["total"],
                           fraction=case["fraction"], history_length=case["history"])
    assert sh.halving_milestones == case["milestones"]
    score = _score_fn(case)
    current = sh.sample()
    for ev, want in enumerate(case["trace"], start=1):
        for c in sorted(current, key=lambda c: (c.alpha, c.seed)):
            sh.update(c, score(c, ev))
        current = sh.sample()
        got = sorted([[c.seed, c.alpha] for c in current])
        assert got == want, f"evaluation {ev}"


def test_state_dict_round_trip_continues_identically():
    case = GOLD["cases"][0]
    pop = _population(case)
    score = _score_fn(case)
    a = SuccessiveHalving(set(pop), case["total"], case["fraction"], case["history"])
    cur = a.sample()
    for ev in range(1, 6):
        for c in sorted(cur, key=lambda c: c.alpha):
            a.update(c, score(c, ev))
        cur = a.sample()
    b = SuccessiveHalving(set(pop), case["total"], case["fraction"], case["history"], state_dict=__import__("pickle").loads(__import__("pickle").dumps(a.state_dict())))  # as tune_alpha.py checkpoints it
    assert b.population == a.population and b.performed_evaluations == a.performed_evaluations
    for ev in range(6, 12):
        for s in (a, b):
            for c in sorted(s.population, key=lambda c: c.alpha):
                s.update(c, score(c, ev))
            s.sample()
        assert a.population == b.population


def test_identity_returns_population_untouched():
    pop = [ExperimentConfig(seed=1, alpha=3.0), ExperimentConfig(seed=2, alpha=30.0)]
    s = Identity(population=pop, total_evaluations=0, state_dict=None)
    s.update(pop[0], 0.5)
    assert s.sample() is pop and s.state_dict() == {"population": pop}
    assert Identity(population=[], total_evaluations=0, state_dict={"population": pop}).sample() is pop
