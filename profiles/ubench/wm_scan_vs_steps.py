"""Debug aid: gradients / parameters of the multistep trainer with the persistent scan kernels vs the per-step launches
(FQLPOP_WM_SCAN is read once per process, hence two child processes) on the inputs of
tests/test_worldmodel_gpu.py::test_multistep_predictor_steps_match_oracle."""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(os.path.dirname(HERE))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, os.path.join(ROOT, "flow-q-learning_b200")); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import test_worldmodel_gpu as tw
    from fqlpop.worldmodel import MULTISTEP, WorldModelTrainer
    T, rows = int(sys.argv[3]), int(sys.argv[4])
    em = tw._envmodel(21)
    em["state"][f"Dense_{len(tw.HID)}"]["kernel"].mul_(0.1)
    tr = WorldModelTrainer(MULTISTEP, tw.O, tw.A, tw.HID, rows, steps=100, params=tw._np(em["state"]),
                           termination_params=tw._np(em["termination"]), termination_hidden_dims=tw.THID, sequence_length=T)
    out = {}
    for step in range(3):
        g = torch.Generator().manual_seed(30 + step)
        obs = torch.randn(rows, T, tw.O, generator=g, dtype=torch.float64)
        batch = {"observations": obs, "actions": torch.rand(rows, T, tw.A, generator=g, dtype=torch.float64) * 2 - 1,
                 "next_observations": obs + 0.05 * torch.randn(rows, T, tw.O, generator=g, dtype=torch.float64),
                 "rewards": -(torch.rand(rows, T, generator=g) >= 0.1).to(torch.float64)}
        tr.step(tw._f32(batch), train=True)
        if step == 0:
            for name, leaf in tw.wmo.tree_leaves(tr.grads()):
                out["g/" + name] = np.asarray(leaf)
    for name, leaf in tw.wmo.tree_leaves(tr.params()):
        out["p/" + name] = np.asarray(leaf)
    np.savez(sys.argv[2], **out)
    sys.exit(0)
for T, rows in ((6, 64), (40, 256)):
    f = {}
    for scan in ("1", "0"):
        path = f"/tmp/wm_{scan}.npz"
        subprocess.run([sys.executable, __file__, "child", path, str(T), str(rows)], check=True, env=dict(os.environ, FQLPOP_WM_SCAN=scan))
        f[scan] = np.load(path)
    print(f"T={T} rows={rows}")
    for k in f["1"].files:
        a, b = f["1"][k].astype(np.float64), f["0"][k].astype(np.float64)
        d = np.abs(a - b)
        print(f"  {k:40s} max|scan - steps| {d.max():.3e}  scale {np.abs(b).max():.3e}  >2e-5: {int((d > 2e-5).sum())} of {d.size}")
