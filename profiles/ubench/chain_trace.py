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
# Euler: 5 layers per step. MMA stamps per layer: start + one per 256-column tile (2,2,2,2,1).
# Epilogue stamps: hidden layer: tile 0 ready, tile 1 ready, act_lo released, act_hi released; head: ready, lo, hi.
mi = ei = 0
for it in range(3):
    for l in range(5):
        nt = 2 if l < 4 else 1
        m = mma[mi:mi + 1 + nt] - t0; mi += 1 + nt
        ne = 4 if l < 4 else 3
        e = epi[ei:ei + ne] - t0; ei += ne
        print(f"it{it} L{l}: mma start {m[0]:7d} tiles issued {[int(x - m[0]) for x in m[1:]]} | epi: tiles ready {[int(x - m[0]) for x in e[:nt]]} released lo/hi {[int(x - m[0]) for x in e[nt:]]}")
