"""A few world-model optimiser steps (state predictor vs frozen termination predictor, antsoccer shape, batch 256, resident
dataset) for an ncu launch list:  ncu --metrics gpu__time_duration.sum --clock-control none --csv python profiles/ubench/wm_steps.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "flow-q-learning_b200"))
from fqlpop.worldmodel import STATE, TERMINATION, WorldModelTrainer, init_params  # noqa: E402

O, A, B, hid, N = 42, 8, 256, (128, 256, 128), 100_000
g = torch.Generator().manual_seed(0)
obs = torch.randn(N, O, generator=g)
data = {"observations": obs, "actions": torch.rand(N, A, generator=g) * 2 - 1,
        "next_observations": obs + 0.05 * torch.randn(N, O, generator=g),
        "rewards": -(torch.rand(N, generator=g) >= 0.005).float()}
tr = WorldModelTrainer(STATE, O, A, hid, B, steps=100, termination_params=init_params(1, TERMINATION, O, A, hid),
                       termination_hidden_dims=hid)
tr.set_dataset(data)
tr.train(int(os.environ.get("WM_STEPS", "8")))
torch.cuda.synchronize()
print(tr.named_logs(True), tr.launches)
