"""Real-environment task (/root/reference/task/offline_task_real.py:12-95).

OUT OF SCOPE of the B200 hot path: it steps ogbench / MuJoCo environments on the CPU. The class keeps the reference's
name and constructor signature so `tune_alpha.py` / `reproduce.py` import it; constructing it needs ogbench, which is
not in this image ("unavailable on this box"). Dataset sampling is shared with the simulated task.
"""
from pathlib import Path

from task.task import Task


class OfflineTaskWithRealEvaluations(Task):
    def __init__(self, env_name: str, buffer_size: int = 2000000, data_directory: Path = None,
                 num_evaluation_envs: int = 50):
        try:
            import ogbench  # noqa: F401
        except ImportError as e:
            raise ImportError(
                "OfflineTaskWithRealEvaluations steps real ogbench/MuJoCo environments, which are not installed here. "
                "Use task.offline_task_simulated.OfflineTaskWithSimulatedEvaluations (learned world model on the GPU)."
            ) from e
        raise NotImplementedError("real-environment evaluation is outside the B200 hot path (SURVEY.md section 2)")

    def sample(self, dataset, batch_size):
        raise NotImplementedError

    def reset(self, seed=None):
        raise NotImplementedError

    def step(self, actions):
        raise NotImplementedError

    def close(self):
        pass
