"""One small simulated evaluation (antsoccer shape, 4 members x 4608 episodes x 40 steps = one wave of 144 CTAs = 72 pairs) for an
`ncu --set full -k regex:fql_rollout --launch-skip 1 --launch-count 1` capture of the fused rollout kernel."""
import os, sys
sys.path.insert(0, "flow-q-learning_b200"); sys.path.insert(0, ".")
import numpy as np, torch
from fqlpop.population import Population
from fqlpop.envmodel import EnvModel
O, A, hidden, P, n, T = 42, 8, (128, 256, 128), 4, 36 * 128, 40
rng = np.random.default_rng(3)
def mlp(i, o, bias_last=0.0, scale_last=1.0):
    dims = [i, *hidden, o]
    t = {f"Dense_{l}": {"kernel": (rng.standard_normal((dims[l], dims[l + 1])) / np.sqrt(dims[l])).astype(np.float32),
                        "bias": np.zeros(dims[l + 1], np.float32)} for l in range(len(dims) - 1)}
    t[f"Dense_{len(dims) - 2}"]["kernel"] *= scale_last
    t[f"Dense_{len(dims) - 2}"]["bias"] += bias_last
    return t
state = mlp(O + A, O, scale_last=0.05)
state["LayerNorm_0"] = {"scale": np.ones(O + A, np.float32), "bias": np.zeros(O + A, np.float32)}
pop = Population(dict(ob_dim=O, action_dim=A, batch_size=256), np.logspace(np.log10(3), 3, P), list(range(P)))
em = EnvModel(state, mlp(O, 1, bias_last=-50.0), O, A)
init = torch.randn(n, O, device="cuda")
for r in range(3):
    out = pop.rollout(em, init, max_episode_steps=T, eval_seed=1 + r)
torch.cuda.synchronize()
print(int(out["steps"].sum()), "lockstep steps")
