"""Dumps the clock64 trace of CTA 0 of the Euler layer-chain kernel (FQLPOP_CHAIN_TRACE=1): per layer, when the MMA
thread started, when it had issued each sub-tile, when the epilogue saw each sub-tile complete and when it released ACT."""
import ctypes, os, sys
os.environ["FQLPOP_CHAIN_TRACE"] = "1"
os.environ["FQLPOP_GROUP_SIZE"] = "0"
sys.path.insert(0, "flow-q-learning_b200"); sys.path.insert(0, ".")
import numpy as np, torch
from fqlpop.population import Population
from fqlpop._lib import lib, check
from bench import synth_dataset
solo = len(sys.argv) > 1 and sys.argv[1] == "solo"
pop = Population(dict(ob_dim=28, action_dim=5, batch_size=256), np.logspace(0.5, 3, 16), list(range(16)))
pop.set_dataset(*synth_dataset(100000, 28, 5, "cuda"))
if solo:
    os.environ["FQLPOP_EAGER"] = "1"
pop.step(3)
torch.cuda.synchronize()
buf = np.zeros(8192, np.int64)
lib.fqlpop_debug_chain_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32]
check(lib.fqlpop_debug_chain_trace(pop._h, buf.ctypes.data, 8192))
mma, epi = buf[:4096], buf[4096:]
t0 = mma[0]
# Euler: 5 layers per step; MMA stamps per layer: 1 + n_sub (4,4,4,4,1); epilogue: hidden: n_sub + 1 (pass-1 end) + 1 (release); head: 1
mi = ei = 0
for it in range(2):
    for l in range(5):
        ns = 4 if l < 4 else 1
        m = mma[mi:mi + 1 + ns] - t0; mi += 1 + ns
        ne = ns + 2 if l < 4 else 1
        e = epi[ei:ei + ne] - t0; ei += ne
        print(f"it{it} L{l}: mma start {m[0]:7d} issued {[int(x - m[0]) for x in m[1:]]}  | epi sub-ready {[int(x - m[0]) for x in e[:ns]] if l < 4 else []} pass1-end/release {[int(x - m[0]) for x in e[ns:]]}")
