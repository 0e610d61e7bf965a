"""Offline replay of a hyper-parameter strategy over logged success curves (reference:
/root/reference/evaluator/strategy_evaluator.py:12-67; driver /root/reference/evaluate_strategy.py).

`evaluate()` walks the strategy exactly like `Trainer.train` does, but every candidate is an `ExperimentFromFile`; it
returns `{step: best score so far}` at the steps where the best score improved (and at the final step). Plotting needs
matplotlib/seaborn, which this image does not have: `plot()` imports them lazily.
"""
from __future__ import annotations

from evaluator.experiment_from_file import ExperimentFromFile


class StrategyEvaluator:
    def __init__(self, strategy, config):
        self.strategy = strategy
        self.config = config
        self.experiments = {}
        self.stopping_step = {}
        self.finished_candidates = set()
        self.candidates = self.strategy.sample()
        for c in self.candidates:
            self.experiments[c] = ExperimentFromFile(experiment_config=c, save_directory=self.config.save_directory,
                                                     env_name=self.config.env_name)

    def evaluate(self) -> dict:
        best_score = 0.0
        score_curve = {}
        current_step = 0
        while True:
            for c in self.candidates:
                if self.experiments[c].train(self.config.eval_interval):
                    self.finished_candidates.add(c)
            current_step += self.config.eval_interval
            score_updated = False
            for c in self.candidates:
                score = self.experiments[c].evaluate(self.config.eval_episodes)
                if score > best_score:
                    best_score = score
                    score_updated = True
                self.strategy.update(c, score)
            if score_updated or current_step == self.config.steps:
                score_curve[current_step] = best_score
            self.candidates = self.strategy.sample()
            if all(c in self.finished_candidates for c in self.candidates):
                break
        return score_curve

    def plot(self):
        import matplotlib.pyplot as plt  # noqa: F401  (optional; absent in this image)
        import seaborn as sns
        sns.set_theme(style="darkgrid")
        plt.figure(figsize=(20, 12))
        for c, exp in self.experiments.items():
            d = exp.get_data()
            sns.lineplot(x=d["step"], y=d["success"], linestyle="--", alpha=0.4)
            upto = [(s, v) for s, v in zip(d["step"], d["success"]) if s <= exp.current_step]
            sns.lineplot(x=[s for s, _ in upto], y=[v for _, v in upto], label=exp.get_label())
        plt.xlabel("step")
        plt.ylabel("success")
        return plt.gcf()
