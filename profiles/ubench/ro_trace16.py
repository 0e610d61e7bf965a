"""Per-layer wait accounting of the whole-rollout kernel (rollout16.cu), CTA (0,0), BASELINE config 3 shape:
MMA issuer: cycles per layer and how many of them it waited for (a) the accumulator to be drained, (b) its A operand from
the epilogue, (c) weight stages from TMA; epilogue threads (lane 0 of the cg = 0 and cg = 3 warps of quarter 0).
Needs the trace build:  make -C flow-q-learning_b200/csrc TRACE=1
Usage: python profiles/ubench/ro_trace16.py [first_layer] [layers]"""
import ctypes, os, sys
HERE = os.path.dirname(os.path.abspath(__file__)); ROOT = os.path.dirname(os.path.dirname(HERE))
os.environ["FQLPOP_LIB"] = os.path.join(ROOT, "flow-q-learning_b200", "lib", "libfqlpop_trace.so")
sys.path.insert(0, os.path.join(ROOT, "flow-q-learning_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
from fqlpop.population import Population
from fqlpop.envmodel import EnvModel
from fqlpop._lib import lib
from bench import synthetic_envmodel_trees
P, n, T, O, A = int(os.environ.get("TRACE_P", "16")), int(os.environ.get("TRACE_N", "4096")), 200, 42, 8
pop = Population(dict(ob_dim=O, action_dim=A, batch_size=256, discount=0.995), np.logspace(0.5, 3, P), list(range(P)))
st, tm = synthetic_envmodel_trees(np, O, A, term_bias=-50.0)
em = EnvModel(st, tm, O, A, device="cuda")
init = torch.randn(n, O, device="cuda")
for _ in range(2):
    pop.rollout(em, init, max_episode_steps=T, eval_seed=1)
torch.cuda.synchronize()
buf = np.zeros(4096, np.int64)
lib.fqlpop_debug_rollout_trace.argtypes = [ctypes.c_void_p, ctypes.c_int32]
assert lib.fqlpop_debug_rollout_trace(buf.ctypes.data, 4096) == 0
mma = buf[:2048].reshape(-1, 8); epi = buf[2048:].reshape(-1, 8)
f0 = int(sys.argv[1]) if len(sys.argv) > 1 else 13
cnt = int(sys.argv[2]) if len(sys.argv) > 2 else 26
print("layer |  MMA: start    total  wait_drained wait_A wait_weights | EPI cg0: enter  full  [bookkept]  done | cg3: enter  full  [bookkept]  done  (rel. to MMA start; full / bookkept: head layers only)")
for l in range(f0, f0 + cnt):
    m, e = mma[l], epi[l]
    if m[0] == 0:
        break
    rel = lambda x: int(x - m[0]) if x else -1
    print(f"{l:5d} | {int(m[0]-mma[f0,0]):10d} {int(m[4]-m[0]):8d} {int(m[1]):9d} {int(m[2]):9d} {int(m[3]):9d}    | "
          f"{rel(e[0]):8d} {rel(e[1]):6d} {rel(e[3]):6d} {rel(e[2]):6d} | {rel(e[4]):8d} {rel(e[5]):6d} {rel(e[7]):6d} {rel(e[6]):6d}")
last = max(i for i in range(len(mma)) if mma[i, 0])
print(f"layers traced {last + 1}; mean cycles per lockstep step {(mma[last, 4] - mma[0, 0]) / (last + 1) * 13:.0f}")
if os.environ.get("TRACE_STAGES"):
    sb = np.zeros(512, np.int64)
    lib.fqlpop_debug_rollout_stage_trace.argtypes = [ctypes.c_void_p]
    assert lib.fqlpop_debug_rollout_stage_trace(sb.ctypes.data) == 0
    peer = sb[256:].reshape(64, 4); sb = sb[:256].reshape(64, 4)
    t0 = sb[0, 0]
    # %globaltimer would be needed to compare the two SMs' clocks exactly; clock64 of the two SMs of a TPC start together
    print("weight stages of step 1, policy layers 1-2 (cycles rel. to the first): slot seen empty | TMA issued | stage seen full | "
          "MMAs issued | full - issued || CTA 1 (pair peer / neighbour tile): slot seen empty | TMA issued")
    for i in range(64):
        a, b, c, d = (int(x - t0) for x in sb[i])
        print(f"{i:3d} {a:8d} {b:8d} {c:8d} {d:8d}   {c - b:6d}   || {int(peer[i, 0] - t0):8d} {int(peer[i, 1] - t0):8d}")
