"""Per-CTA timeline of the layer-chain launches of one graph-replayed population step (FQLPOP_TIMELINE=1): for every
chain launch (tag = launch index within the step) the first start / last end relative to the step's first CTA, the
CTA count and the mean CTA duration; then SM occupancy by chain CTAs over the step.
Usage: python profiles/ubench/timeline.py [members] [group_size]"""
import ctypes, os, sys
os.environ["FQLPOP_TIMELINE"] = "1"
if len(sys.argv) > 2:
    os.environ["FQLPOP_GROUP_SIZE"] = sys.argv[2]
sys.path.insert(0, "flow-q-learning_b200"); sys.path.insert(0, ".")
import numpy as np, torch
from fqlpop.population import Population
from fqlpop._lib import lib, check
from bench import synth_dataset
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
pop = Population(dict(ob_dim=28, action_dim=5, batch_size=256), np.logspace(0.5, 3, n), list(range(n)))
pop.set_dataset(*synth_dataset(1000000, 28, 5, "cuda"))
lib.fqlpop_debug_timeline.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p]
buf = np.zeros((65536, 4), np.int64)
cnt = ctypes.c_int32(0)
pop.step(20)
torch.cuda.synchronize()
check(lib.fqlpop_debug_timeline(pop._h, buf.ctypes.data, 65536, ctypes.byref(cnt)))
NS = int(os.environ.get('TL_STEPS', '1'))
pop.step(NS)
torch.cuda.synchronize()
check(lib.fqlpop_debug_timeline(pop._h, buf.ctypes.data, 65536, ctypes.byref(cnt)))
r = buf[:cnt.value]
if NS > 1:  # keep the last step of every (group, launch): its CTAs are the last ones in time
    keep = []
    for key in set(r[:, 3] & 0xFFFFFF):
        m = np.where((r[:, 3] & 0xFFFFFF) == key)[0]
        m = m[np.argsort(r[m, 0])]
        keep.extend(m[-(len(m) // NS):])
    r = r[np.array(keep)]
t0 = r[:, 0].min()
names = {0: "euler", 1: "pi_fwd", 2: "tq_fwd", 3: "qc_fwd", 4: "qc_bwd", 5: "bc_fwd", 6: "bc_bwd", 7: "qa_fwd", 8: "qa_bwd", 9: "pi_bwd",
         20: "pi_head", 21: "tq_head", 22: "qc_head", 23: "qa_head", 24: "bc_head", 30: "adam_q", 31: "adam_bc", 32: "adam_pi"}
print(f"records {len(r)}  step span {(r[:, 1].max() - t0) / 1e3:.1f} us")
tag = r[:, 3] & 0xFF
grp = (r[:, 3] >> 8) & 0xFFFF
for g in sorted(set(grp)):
    for k in sorted(names):
        m = (tag == k) & (grp == g)
        if not m.any():
            continue
        s, e = r[m, 0], r[m, 1]
        print(f"group@{g:3d} {names[k]:7s} ctas {m.sum():3d} start {(s.min() - t0) / 1e3:7.1f} .. {(s.max() - t0) / 1e3:7.1f}"
              f"  end {(e.max() - t0) / 1e3:7.1f}  mean cta {np.mean(e - s) / 1e3:6.1f} us")
r = r[tag < 10]
busy = (r[:, 1] - r[:, 0]).sum() / 1e3
print(f"chain CTA-us total {busy:.0f}; / 148 SMs = {busy / 148:.1f} us; SMs used {len(set(r[:, 2]))}")

# ---- occupancy profile (TL_PROFILE=bin_us): chain / adam / head CTAs resident per time bin over the recorded window
if os.environ.get("TL_PROFILE"):
    binus = float(os.environ["TL_PROFILE"])
    r = buf[:cnt.value]
    t0 = r[:, 0].min(); span = (r[:, 1].max() - t0) / 1e3
    nb = int(span / binus) + 1
    tag = r[:, 3] & 0xFF
    kinds = {"chain": tag < 10, "head": (tag >= 20) & (tag < 30), "adam": tag >= 30}
    prof = {k: np.zeros(nb) for k in kinds}
    for k, m in kinds.items():
        for s, e in zip((r[m, 0] - t0) / 1e3, (r[m, 1] - t0) / 1e3):
            b0, b1 = int(s / binus), int(e / binus)
            for b in range(b0, b1 + 1):
                lo, hi = max(s, b * binus), min(e, (b + 1) * binus)
                if hi > lo:
                    prof[k][b] += (hi - lo) / binus
    print(f"occupancy profile, {binus} us bins, mean resident CTAs (chain CTAs fill an SM; adam CTAs fill an SM; heads do not)")
    for b in range(nb):
        print(f"{b * binus:8.0f} us  chain {prof['chain'][b]:6.1f}  adam {prof['adam'][b]:6.1f}  head {prof['head'][b]:6.1f}")
    print("totals CTA-us:", {k: float(v.sum() * binus) for k, v in prof.items()}, "span", span)
