"""Clock stamps of the traced epilogue thread (CTA 0, first epilogue warp) and of the MMA issuer of one layer-chain
launch of the step (FQLPOP_TRACE_LAUNCH: 0 euler, 1 pi_fwd, 2 tq_fwd, 3 qc_fwd, 4 qc_bwd, 5 bc_fwd, 6 bc_bwd,
7 qa_fwd, 8 qa_bwd, 9 pi_bwd). Usage: python profiles/ubench/epi_trace.py <launch> [max_stamps]"""
import ctypes, os, sys
os.environ["FQLPOP_CHAIN_TRACE"] = "1"
os.environ["FQLPOP_TRACE_LAUNCH"] = sys.argv[1] if len(sys.argv) > 1 else "0"
sys.path.insert(0, "flow-q-learning_b200"); sys.path.insert(0, ".")
import numpy as np, torch
from fqlpop.population import Population
from fqlpop._lib import lib, check
from bench import synth_dataset
pop = Population(dict(ob_dim=28, action_dim=5, batch_size=256), np.logspace(0.5, 3, 16), list(range(16)))
pop.set_dataset(*synth_dataset(100000, 28, 5, "cuda"))
pop.step(5)
torch.cuda.synchronize()
buf = np.zeros(8192, np.int64)
lib.fqlpop_debug_chain_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32]
check(lib.fqlpop_debug_chain_trace(pop._h, buf.ctypes.data, 8192))
mma = buf[:4096]
n = int(buf[4096])
st = buf[4098:4098 + 2 * n].reshape(-1, 2)
t0 = min(mma[0], st[0, 0]) if n else mma[0]
print("MMA issuer stamps (cycles from t0):", [int(x - t0) for x in mma[:int(os.environ.get("NMMA", "24"))] if x > 0])
lim = int(sys.argv[2]) if len(sys.argv) > 2 else 120
prev = t0
for c, i in st[:lim]:
    print(f"  id {int(i):3d}  t {int(c - t0):8d}  +{int(c - prev):7d}")
    prev = c
