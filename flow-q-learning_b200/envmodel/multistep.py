"""MultistepStatePredictor descriptor (constructor arguments of /root/reference/envmodel/multistep.py:34-37): the baseline
cell scanned over a window of `sequence_length` actions, each prediction fed back as the next observation. Training runs
in libfqlpop.so (fqlpop_wm_* with kind 2: back-propagation through time); the checkpoint keeps the reference's nesting
params["ScanCell_0"]["cell"], which utils/envmodel.load_model unwraps for the rollout."""
from dataclasses import dataclass
from typing import Tuple


@dataclass(frozen=True)
class MultistepStatePredictor:
    observation_dimension: int = 28
    action_dimension: int = 5
    hidden_dims: Tuple[int, ...] = (128, 256, 128)
