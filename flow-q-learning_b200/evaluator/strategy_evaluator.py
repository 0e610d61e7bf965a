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

    def _round(self, step: int, best: float):
        """One strategy round at `step`: advance every live candidate's log cursor, score it, feed the strategy.
        Returns the best score after the round."""
        interval, episodes = self.config.eval_interval, self.config.eval_episodes
        self.finished_candidates.update(c for c in self.candidates if self.experiments[c].train(interval))
        for c in self.candidates:  # strategy.update order = candidate order, as in Trainer.train
            score = self.experiments[c].evaluate(episodes)
            best = max(best, score)
            self.strategy.update(c, score)
        return best

    def evaluate(self) -> dict:
        curve, best, step = {}, 0.0, 0
        live = True
        while live:
            step += self.config.eval_interval
            new_best = self._round(step, best)
            if new_best > best or step == self.config.steps:  # a point where the incumbent improved, and the last step
                curve[step] = new_best
            best = new_best
            self.candidates = self.strategy.sample()
            live = not self.finished_candidates.issuperset(self.candidates)
        return curve

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
