"""How much of a layer-chain launch is the weight stream? Times the Euler chain (which=0) and the one-step policy pass
(which=1) of 16 members back to back, normally and with FQLPOP_CHAIN_DEBUG=1 (no weight TMA, no stage waits: the bare
MMA + epilogue stream; results are garbage). Run twice:  python profiles/ubench/weight_stream.py ; FQLPOP_CHAIN_DEBUG=1 python ..."""
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
s = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for which, name, layers in ((0, "euler chain (10 steps x 5 layers, 32 CTAs)", 50), (1, "policy pass (4 hidden layers, 96 CTAs)", 4)):
    check(lib.fqlpop_debug_chain_launch(pop._h, which, 3, s))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(lib.fqlpop_debug_chain_launch(pop._h, which, 20, s))
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1e3
    print("FQLPOP_CHAIN_DEBUG=%s %s: %.1f us per launch, %.2f us per layer" % (os.environ.get("FQLPOP_CHAIN_DEBUG", "0"), name, us, us / layers))
