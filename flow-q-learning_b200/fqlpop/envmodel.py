"""Device-resident learned world model (state predictor + termination predictor) and the fused rollout call.

Host-side mirror of what the reference reaches through `load_model(...) -> apply_fn`
(/root/reference/utils/envmodel.py:9-56) and `OfflineTaskWithSimulatedEvaluations.step`
(/root/reference/task/offline_task_simulated.py:85-107). Parameters arrive as flax-style trees
(`LayerNorm_0/{scale,bias}`, `Dense_i/{kernel[in,out],bias}`); all arithmetic runs in libfqlpop.so.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional

import numpy as np
import torch

from . import _lib
from ._lib import check, lib

_vp = ctypes.c_void_p


class EnvModelConfig(ctypes.Structure):
    _fields_ = [
        ("ob_dim", ctypes.c_int32), ("action_dim", ctypes.c_int32),
        ("n_state_hidden", ctypes.c_int32), ("state_hidden", ctypes.c_int32 * 6),
        ("n_term_hidden", ctypes.c_int32), ("term_hidden", ctypes.c_int32 * 6),
        ("ln_eps", ctypes.c_float),
    ]


class RolloutArgs(ctypes.Structure):
    _fields_ = [
        ("n_episodes", ctypes.c_int32), ("max_episode_steps", ctypes.c_int32), ("horizon", ctypes.c_int32),
        ("init_obs", _vp), ("init_obs_member_stride", ctypes.c_int64),
        ("noise", _vp), ("noise_member_stride", ctypes.c_int64),
        ("eval_seed", ctypes.c_uint64),
        ("out_success", _vp), ("out_length", _vp), ("out_success_rate", _vp), ("out_return_mean", _vp),
        ("out_steps", _vp), ("traj_obs", _vp), ("traj_act", _vp), ("traj_logit", _vp),
    ]


lib.fqlpop_envmodel_create.argtypes = [ctypes.POINTER(EnvModelConfig), _vp, _vp, _vp, _vp, _vp, _vp,
                                       ctypes.POINTER(_vp)]
lib.fqlpop_envmodel_destroy.argtypes = [_vp]
lib.fqlpop_rollout.argtypes = [_vp, _vp, ctypes.POINTER(RolloutArgs), _vp]
lib.fqlpop_sample_actions.argtypes = [_vp, ctypes.c_int32, _vp, ctypes.c_int32, _vp, ctypes.c_uint64, _vp, _vp]
lib.fqlpop_env_step.argtypes = [_vp, _vp, _vp, ctypes.c_int32, _vp, _vp, _vp]
lib.fqlpop_rollout_launches.argtypes = [_vp]
lib.fqlpop_rollout_launches.restype = ctypes.c_int64
for _f in ("envmodel_create", "envmodel_destroy", "rollout", "sample_actions", "env_step"):
    getattr(lib, "fqlpop_" + _f).restype = ctypes.c_int


def _dense_layers(tree) -> list:
    n = sum(1 for k in tree if k.startswith("Dense_"))
    return [tree[f"Dense_{i}"] for i in range(n)]


def _f32(x) -> np.ndarray:
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.ascontiguousarray(np.asarray(x, dtype=np.float32))


class EnvModel:
    """state: tree of BaselineStatePredictor (envmodel/baseline.py:25-37; for the `multistep` model the
    reference unwraps params["params"]["ScanCell_0"]["cell"], utils/envmodel.py:46-49 -- pass that subtree);
    termination: tree of TerminationPredictor (envmodel/termination_predictor.py:14-21)."""

    def __init__(self, state: Dict, termination: Dict, ob_dim: int, action_dim: int, ln_eps: float = 1e-6,
                 device=None):
        if not torch.cuda.is_available():
            raise _lib.FqlPopError("no CUDA device: the world-model rollout has no CPU fallback")
        self.device = torch.device(device) if device is not None else _lib.default_device()
        if self.device.index is None:   # a bare "cuda" means the current device
            self.device = torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        sl, tl = _dense_layers(state), _dense_layers(termination)
        if len(sl) < 2 or len(tl) < 2:
            raise ValueError("world model networks need at least one hidden layer")
        c = EnvModelConfig(ob_dim=int(ob_dim), action_dim=int(action_dim), n_state_hidden=len(sl) - 1,
                           n_term_hidden=len(tl) - 1, ln_eps=float(ln_eps))
        sk, sb = [_f32(d["kernel"]) for d in sl], [_f32(d["bias"]) for d in sl]
        tk, tb = [_f32(d["kernel"]) for d in tl], [_f32(d["bias"]) for d in tl]
        k_in = ob_dim + action_dim
        for i, k in enumerate(sk):
            assert k.shape[0] == k_in, f"state predictor Dense_{i}: kernel {k.shape}, expected in={k_in}"
            k_in = k.shape[1]
            if i < len(sk) - 1:
                c.state_hidden[i] = k.shape[1]
        assert k_in == ob_dim, "state predictor output must be ob_dim wide"
        k_in = ob_dim
        for i, k in enumerate(tk):
            assert k.shape[0] == k_in, f"termination predictor Dense_{i}: kernel {k.shape}, expected in={k_in}"
            k_in = k.shape[1]
            if i < len(tk) - 1:
                c.term_hidden[i] = k.shape[1]
        assert k_in == 1, "termination predictor output must be a single logit"
        ln_s, ln_b = _f32(state["LayerNorm_0"]["scale"]), _f32(state["LayerNorm_0"]["bias"])
        assert ln_s.size == ob_dim + action_dim

        def ptrs(arrs):
            return (ctypes.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])

        h = _vp()
        check(lib.fqlpop_envmodel_create(ctypes.byref(c), ptrs(sk), ptrs(sb), ln_s.ctypes.data, ln_b.ctypes.data,
                                         ptrs(tk), ptrs(tb), ctypes.byref(h)), "fqlpop_envmodel_create")
        self._h = h
        self.O, self.A = int(ob_dim), int(action_dim)

    def close(self):
        if getattr(self, "_h", None):
            lib.fqlpop_envmodel_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def step(self, observations, actions):
        """Task.step arithmetic for a batch: returns (next_observations [n,O], termination logits [n]) on device."""
        obs = torch.as_tensor(observations, dtype=torch.float32).to(self.device).contiguous()
        act = torch.as_tensor(actions, dtype=torch.float32).to(self.device).contiguous()
        n = obs.shape[0]
        assert obs.shape == (n, self.O) and act.shape == (n, self.A)
        nxt = torch.empty_like(obs)
        logits = torch.empty(n, dtype=torch.float32, device=self.device)
        s = torch.cuda.current_stream(self.device).cuda_stream
        check(lib.fqlpop_env_step(self._h, obs.data_ptr(), act.data_ptr(), n, nxt.data_ptr(), logits.data_ptr(), s),
              "fqlpop_env_step")
        return nxt, logits


def rollout(pop, envmodel: EnvModel, init_obs, max_episode_steps: int, horizon: int = 0, noise=None,
            eval_seed: int = 0, trajectory: bool = False, stream: Optional[torch.cuda.Stream] = None):
    """One fused simulated evaluation of every active member of `pop`.

    init_obs: [n,O] (shared by all members) or [P,n,O]; noise: None (device Philox) or [T,n,A] / [P,T,n,A].
    Returns a dict of device tensors indexed by member id: success [P,n] uint8, length [P,n] int32,
    success_rate [P], return_mean [P], steps [P] (+ traj_obs/traj_act/traj_logit when trajectory=True).
    """
    dev = pop.device
    P, O, A = pop.P, pop.O, pop.A
    io = torch.as_tensor(init_obs, dtype=torch.float32).to(dev).contiguous()
    shared = io.dim() == 2
    n = io.shape[-2]
    assert io.shape[-1] == O and (shared or io.shape[0] == P)
    T = horizon if 0 < horizon < max_episode_steps else max_episode_steps
    nz = None
    nz_stride = 0
    if noise is not None:
        nz = torch.as_tensor(noise, dtype=torch.float32).to(dev).contiguous()
        assert nz.shape[-3:] == (T, n, A), (tuple(nz.shape), (T, n, A))
        nz_stride = 0 if nz.dim() == 3 else T * n * A
    out = {
        "success": torch.zeros(P, n, dtype=torch.uint8, device=dev),
        "length": torch.zeros(P, n, dtype=torch.int32, device=dev),
        "success_rate": torch.zeros(P, dtype=torch.float32, device=dev),
        "return_mean": torch.zeros(P, dtype=torch.float32, device=dev),
        "steps": torch.zeros(P, dtype=torch.int32, device=dev),
    }
    if trajectory:
        out["traj_obs"] = torch.zeros(P, T, n, O, dtype=torch.float32, device=dev)
        out["traj_act"] = torch.zeros(P, T, n, A, dtype=torch.float32, device=dev)
        out["traj_logit"] = torch.zeros(P, T, n, dtype=torch.float32, device=dev)
    a = RolloutArgs(
        n_episodes=n, max_episode_steps=int(max_episode_steps), horizon=int(horizon),
        init_obs=io.data_ptr(), init_obs_member_stride=0 if shared else n * O,
        noise=None if nz is None else nz.data_ptr(), noise_member_stride=nz_stride, eval_seed=int(eval_seed),
        out_success=out["success"].data_ptr(), out_length=out["length"].data_ptr(),
        out_success_rate=out["success_rate"].data_ptr(), out_return_mean=out["return_mean"].data_ptr(),
        out_steps=out["steps"].data_ptr(),
        traj_obs=out["traj_obs"].data_ptr() if trajectory else None,
        traj_act=out["traj_act"].data_ptr() if trajectory else None,
        traj_logit=out["traj_logit"].data_ptr() if trajectory else None,
    )
    s = (stream or torch.cuda.current_stream(dev)).cuda_stream
    check(lib.fqlpop_rollout(pop._h, envmodel._h, ctypes.byref(a), s), "fqlpop_rollout")
    out["_keep"] = (io, nz)
    return out


def sample_actions(pop, member: int, observations, noise=None, seed: int = 0,
                   stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
    """FQLAgent.sample_actions for one member: clip(pi(obs, z), -1, 1) -> [n, A] device tensor."""
    dev = pop.device
    obs = torch.as_tensor(observations, dtype=torch.float32).to(dev).contiguous()
    if obs.dim() == 1:
        obs = obs.unsqueeze(0)
    n = obs.shape[0]
    assert obs.shape == (n, pop.O)
    nz = None if noise is None else torch.as_tensor(noise, dtype=torch.float32).to(dev).contiguous()
    if nz is not None:
        assert nz.shape == (n, pop.A)
    out = torch.empty(n, pop.A, dtype=torch.float32, device=dev)
    s = (stream or torch.cuda.current_stream(dev)).cuda_stream
    check(lib.fqlpop_sample_actions(pop._h, int(member), obs.data_ptr(), n, None if nz is None else nz.data_ptr(),
                                    int(seed) & 0xFFFFFFFFFFFFFFFF, out.data_ptr(), s), "fqlpop_sample_actions")
    return out
