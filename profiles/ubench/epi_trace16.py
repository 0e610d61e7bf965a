"""Per-layer wait accounting of one chain launch of the step, second-generation kernel (chain16.cu), CTA 0:
MMA issuer: cycles per layer and how many of them it waited for (a) the accumulator to be drained, (b) ACT blocks from
the epilogue, (c) weight stages from TMA; epilogue thread (first epilogue warp, lane 0): stamps at tile 0 ready / pass 0
done / tile 1 ready / block cg published / layer done / copies done.
Needs the trace build:  make -C flow-q-learning_b200/csrc TRACE=1
Usage: python profiles/ubench/epi_trace16.py <launch> [layers]   launch: 0 euler, 1 pi_fwd, 2 tq_fwd, 3 qc_fwd, 4 qc_bwd,
5 bc_fwd, 6 bc_bwd, 7 qa_fwd, 8 qa_bwd, 9 pi_bwd"""
import ctypes, os, sys
HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(os.path.dirname(HERE))
os.environ["FQLPOP_LIB"] = os.path.join(ROOT, "flow-q-learning_b200", "lib", "libfqlpop_trace.so")
os.environ["FQLPOP_CHAIN_TRACE"] = "1"
os.environ["FQLPOP_TRACE_LAUNCH"] = sys.argv[1] if len(sys.argv) > 1 else "0"
sys.path.insert(0, os.path.join(ROOT, "flow-q-learning_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
from fqlpop.population import Population
from fqlpop._lib import lib, check
from bench import synth_dataset
P = int(os.environ.get("TRACE_P", "16"))
pop = Population(dict(ob_dim=28, action_dim=5, batch_size=256), np.logspace(0.5, 3, P), list(range(P)))
pop.set_dataset(*synth_dataset(1000000, 28, 5, "cuda"))
pop.step(10)
torch.cuda.synchronize()
buf = np.zeros(8192, np.int64)
lib.fqlpop_debug_chain_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32]
check(lib.fqlpop_debug_chain_trace(pop._h, buf.ctypes.data, 8192))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 12
mma = buf[:4096].reshape(-1, 8); epi = buf[4096:].reshape(-1, 8)
t0 = mma[0, 0]
print(f"launch {os.environ['FQLPOP_TRACE_LAUNCH']}  (cycles; CTA 0)")
print("layer |  MMA: start    total  wait_drained wait_act wait_weights | EPI: t0rdy  pass0  t1rdy  pub0   end  copied (rel. to MMA start)")
for l in range(n):
    m, e = mma[l], epi[l]
    if m[0] == 0:
        break
    rel = lambda x: int(x - m[0]) if x else -1
    print(f"{l:5d} | {int(m[0]-t0):10d} {int(m[4]-m[0]):8d} {int(m[1]):9d} {int(m[2]):9d} {int(m[3]):9d}    | "
          f"{rel(e[1]):8d} {rel(e[2]):6d} {rel(e[3]):6d} {rel(e[4]):6d} {rel(e[5]):6d} {rel(e[6]):6d}")
last = max(i for i in range(len(mma)) if mma[i, 0])
print(f"layers traced {last + 1}; mean cycles per layer {(mma[last, 4] - t0) / (last + 1):.0f}")
