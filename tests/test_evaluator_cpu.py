"""ExperimentFromFile / StrategyEvaluator against golden score curves produced by EXECUTING the reference's
/root/reference/evaluator/{experiment_from_file,strategy_evaluator}.py (tests/golden/make_strategy_evaluator.py)."""
import json
import os
from dataclasses import dataclass
from pathlib import Path

import pytest

from evaluator.experiment_from_file import ExperimentFromFile, config_to_name
from evaluator.strategy_evaluator import StrategyEvaluator
from hpo.successive_halving import SuccessiveHalving
from trainer.config import ExperimentConfig

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "strategy_evaluator.json")))


@dataclass
class _Cfg:
    steps: int
    eval_interval: int
    save_directory: Path
    env_name: str = "cube-single-play-singletask-task2-v0"
    eval_episodes: int = 50


def _write_runs(root, env, configs, curves, interval):
    for c, cur in zip(configs, curves):
        # an older run directory with different numbers: the newest timestamp must win
        for stamp, scale in (("20240101_000000", 0.0), ("20250607_101112", 1.0)):
            d = root / env / f"{config_to_name(c)}_{stamp}"
            d.mkdir(parents=True)
            with open(d / "eval.csv", "w") as f:
                f.write("success,step\n")
                for k, v in enumerate(cur):
                    f.write(f"{v * scale},{(k + 1) * interval}\n")


@pytest.mark.parametrize("case", GOLD["cases"], ids=lambda c: f"{c['kind']}-n{c['n']}-f{c['fraction']}-h{c['history_length']}")
def test_strategy_replay_matches_reference(case, tmp_path):
    configs = [ExperimentConfig(seed=int(s), alpha=a) for s, a in case["configs"]]
    assert [config_to_name(c) for c in configs] == case["names"]
    cfg = _Cfg(steps=case["steps"], eval_interval=case["eval_interval"], save_directory=tmp_path)
    _write_runs(tmp_path, cfg.env_name, configs, case["curves"], case["eval_interval"])
    strat = SuccessiveHalving(population=set(configs), total_evaluations=case["total_evaluations"],
                              fraction=case["fraction"], history_length=case["history_length"])
    ev = StrategyEvaluator(strat, cfg)
    curve = ev.evaluate()
    # same steps; values to 1e-14: the reference reads eval.csv with pandas' default (fast, not round-trip) float parser,
    # which is off by one ulp on some of these numbers, this build parses them exactly
    got = {str(k): v for k, v in curve.items()}
    assert list(got) == list(case["score_curve"])
    assert list(got.values()) == pytest.approx(list(case["score_curve"].values()), rel=1e-14, abs=0)
    assert {config_to_name(c): e.current_step for c, e in ev.experiments.items()} == case["stopped_at"]


def test_experiment_from_file_edges(tmp_path):
    c = ExperimentConfig(seed=3, alpha=10.0)
    assert config_to_name(c) == "seed_3_alpha_10.0"
    assert config_to_name(ExperimentConfig(alpha=0.5)) == "alpha_0.5"            # None fields are skipped
    with pytest.raises(FileNotFoundError):
        ExperimentFromFile(c, tmp_path, "env")
    _write_runs(tmp_path, "env", [c], [[0.1, 0.2, 0.3]], 100)
    e = ExperimentFromFile(c, tmp_path, "env")
    assert e.max_steps == 300 and e.current_step == 0
    assert e.train(100) is False and e.evaluate() == 0.1
    assert e.train(100) is False and e.evaluate() == 0.2
    assert e.train(100) is True and e.evaluate() == 0.3
    assert e.train(100) is True
    with pytest.raises(ValueError):                                               # no row for step 400
        e.evaluate()
    assert e.get_data() == {"step": [100, 200, 300], "success": [0.1, 0.2, 0.3]}
    assert "seed = 3" in e.get_label() and "10.00" in e.get_label()
