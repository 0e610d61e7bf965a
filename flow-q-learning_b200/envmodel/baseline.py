"""Network descriptors with the reference's constructor arguments. The arithmetic is in libfqlpop.so (wm.cu for training,
rollout.cu for evaluation); these objects only carry the shapes the trainers need.
BaselineStatePredictor: /root/reference/envmodel/baseline.py:17-37 (LayerNorm(concat(o, a)) -> [Dense, relu]* -> Dense(O) + o).
TerminationPredictor:   /root/reference/envmodel/termination_predictor.py:8-21 (Dropout(0.1) -> [Dense, relu]* -> Dense(1))."""
from dataclasses import dataclass
from typing import Tuple


@dataclass(frozen=True)
class BaselineStatePredictor:
    observation_dimension: int = 28
    action_dimension: int = 5
    hidden_dims: Tuple[int, ...] = (128, 128)


@dataclass(frozen=True)
class TerminationPredictor:
    observation_dimension: int = 28
    hidden_dims: Tuple[int, ...] = (128, 128)
    dropout_rate: float = 0.1
