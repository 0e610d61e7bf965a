"""Times optimiser steps of the multistep world-model trainer (envmodel/multistep.py: the cell scanned over a window of
T = 256 steps, batch 256, antsoccer shape, frozen termination predictor, resident dataset of whole 1000-step episodes).
FQLPOP_WM_SCAN=0 selects the per-step launches instead of the persistent scan kernels."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "flow-q-learning_b200"))
from fqlpop.worldmodel import MULTISTEP, TERMINATION, WorldModelTrainer, init_params  # noqa: E402

O, A, B, T, hid, N = 42, 8, 256, int(os.environ.get("WM_T", "256")), (128, 256, 128), 200_000
tw = float(os.environ.get("WM_TERM_WEIGHT", "1.0"))
g = torch.Generator().manual_seed(0)
obs = torch.randn(N, O, generator=g)
data = {"observations": obs, "actions": torch.rand(N, A, generator=g) * 2 - 1,
        "next_observations": obs + 0.05 * torch.randn(N, O, generator=g),
        "rewards": -(torch.rand(N, generator=g) >= 0.005).float()}
kw = dict(termination_params=init_params(1, TERMINATION, O, A, hid), termination_hidden_dims=hid) if tw > 0 else {}
tr = WorldModelTrainer(MULTISTEP, O, A, hid, B, steps=10_000, sequence_length=T, termination_weight=tw, **kw)
tr.set_dataset(data)
tr.train(3)
torch.cuda.synchronize()
l0 = tr.launches
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = int(os.environ.get("WM_STEPS", "20"))
e0.record()
tr.train(n)
e1.record()
torch.cuda.synchronize()
print(f"scan={os.environ.get('FQLPOP_WM_SCAN', '1')} T={T} termination_weight={tw}: {e0.elapsed_time(e1) / n:.3f} ms per optimiser step, "
      f"{(tr.launches - l0) // n} launches per step, loss {tr.named_logs(True).get('loss')}")
