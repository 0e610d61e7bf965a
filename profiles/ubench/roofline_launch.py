"""The launch bench.py's `roofline` block times (one-step policy forward pass of all 16 members, 96 CTAs of
fql_chain_kernel<0>), run 3 times after 3 population steps -- for `ncu --set full -k regex:fql_chain_kernel
--launch-skip 120 --launch-count 3` (3 steps x 4 groups x 10 chain launches precede it)."""
import ctypes, os, sys
sys.path.insert(0, "flow-q-learning_b200"); sys.path.insert(0, ".")
import numpy as np, torch
from fqlpop.population import Population
from fqlpop._lib import lib, check
from bench import synth_dataset
pop = Population(dict(ob_dim=28, action_dim=5, batch_size=256), np.logspace(np.log10(3), 3, 16), list(range(16)))
pop.set_dataset(*synth_dataset(1000000, 28, 5, "cuda"))
pop.step(3)
torch.cuda.synchronize()
lib.fqlpop_debug_chain_launch.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
which = int(sys.argv[1]) if len(sys.argv) > 1 else 1   # 0 euler, 1 policy fwd, 2 critic fwd (LN+save), 3 critic bwd, 4 target critic
check(lib.fqlpop_debug_chain_launch(pop._h, which, 3, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
torch.cuda.synchronize()
print("done")
