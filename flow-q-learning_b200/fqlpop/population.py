"""ctypes binding of the population entry points of libfqlpop.so (include/fqlpop.h).

`Population` owns a device-resident population of FQL agents. It is the object underneath the reference-facing
`FQLAgent` views (fqlpop.agent) and the lockstep `Trainer` (trainer/trainer.py). torch is used only to hold device
buffers and streams; all arithmetic happens in the CUDA library. There is no CPU fallback.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import _lib
from ._lib import check, lib

INFO_KEYS = (
    "critic/critic_loss", "critic/q_mean", "critic/q_max", "critic/q_min",
    "actor/actor_loss", "actor/bc_flow_loss", "actor/distill_loss", "actor/q_loss", "actor/q", "actor/mse",
    "grad/max", "grad/min", "grad/norm",
)
INFO_SLOTS = 16
NETS = {"critic": 0, "actor_bc_flow": 1, "actor_onestep_flow": 2}
WHICH = {"params": 0, "mu": 1, "nu": 2, "target": 3, "grad": 4}


class Config(ctypes.Structure):
    _fields_ = [
        ("ob_dim", ctypes.c_int32), ("action_dim", ctypes.c_int32),
        ("hidden_actor", ctypes.c_int32), ("hidden_value", ctypes.c_int32),
        ("n_hidden_actor", ctypes.c_int32), ("n_hidden_value", ctypes.c_int32),
        ("num_ensembles", ctypes.c_int32),
        ("flow_steps", ctypes.c_int32), ("batch_size", ctypes.c_int32), ("population", ctypes.c_int32),
        ("lr", ctypes.c_float), ("discount", ctypes.c_float), ("tau", ctypes.c_float),
        ("adam_b1", ctypes.c_float), ("adam_b2", ctypes.c_float), ("adam_eps", ctypes.c_float),
        ("ln_eps", ctypes.c_float),
        ("q_agg_min", ctypes.c_int32), ("normalize_q_loss", ctypes.c_int32),
        ("layer_norm", ctypes.c_int32), ("actor_layer_norm", ctypes.c_int32),
    ]


_vp = ctypes.c_void_p
lib.fqlpop_create.argtypes = [ctypes.POINTER(Config), _vp, _vp, ctypes.POINTER(_vp)]
lib.fqlpop_destroy.argtypes = [_vp]
lib.fqlpop_set_dataset.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp, ctypes.c_int64]
lib.fqlpop_set_noise.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp, _vp]
lib.fqlpop_set_active.argtypes = [_vp, _vp, ctypes.c_int32]
lib.fqlpop_step.argtypes = [_vp, ctypes.c_int32, _vp]
lib.fqlpop_total_loss.argtypes = [_vp, _vp]
lib.fqlpop_info.argtypes = [_vp, _vp, ctypes.c_int32, _vp]
lib.fqlpop_info_async.argtypes = [_vp, _vp, _vp]
lib.fqlpop_leaf.argtypes = [_vp] + [ctypes.c_int32] * 5 + [_vp, ctypes.c_int32]
lib.fqlpop_refresh_shadow.argtypes = [_vp, _vp]
lib.fqlpop_counters.argtypes = [_vp, _vp, _vp, ctypes.c_int32]
lib.fqlpop_launches_per_step.argtypes = [_vp]
lib.fqlpop_last_batch.argtypes = [_vp, _vp, _vp]
lib.fqlpop_last_batch.restype = ctypes.c_int
lib.fqlpop_philox_fill.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                   ctypes.c_uint32, _vp]
lib.fqlpop_philox_fill.restype = ctypes.c_int
for _f in ("create", "destroy", "set_dataset", "set_noise", "set_active", "step", "total_loss", "info", "leaf",
           "refresh_shadow", "counters", "launches_per_step", "member_export", "member_import"):
    getattr(lib, "fqlpop_" + _f).restype = ctypes.c_int
lib.fqlpop_member_state_words.argtypes = [_vp]
lib.fqlpop_member_state_words.restype = ctypes.c_int64
lib.fqlpop_member_export.argtypes = [_vp, ctypes.c_int32, _vp, _vp]
lib.fqlpop_member_import.argtypes = [_vp, ctypes.c_int32, _vp, _vp]


def _uniform_width(dims, what):
    dims = tuple(int(d) for d in dims)
    if len(set(dims)) != 1:
        raise _lib.FqlPopError(f"{what}: only uniform hidden widths are supported, got {dims}")
    return dims[0], len(dims)


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


class Population:
    """P lockstep FQL agents on one GPU. `agent_config` is any mapping/object with the AgentConfig fields
    (/root/reference/trainer/config.py:6-23) plus ob_dim / action_dim."""

    def __init__(self, agent_config, alphas: Sequence[float], seeds: Sequence[int], device=None):
        get = (lambda k, d=None: agent_config.get(k, d)) if isinstance(agent_config, dict) else (
            lambda k, d=None: getattr(agent_config, k, d))
        if not torch.cuda.is_available():
            raise _lib.FqlPopError("no CUDA device: the population hot path has no CPU fallback")
        self.device = torch.device(device) if device is not None else _lib.default_device()
        if self.device.index is None:   # a bare "cuda" means the current device
            self.device = torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        ha, nha = _uniform_width(get("actor_hidden_dims", (512,) * 4), "actor_hidden_dims")
        hv, nhv = _uniform_width(get("value_hidden_dims", (512,) * 4), "value_hidden_dims")
        ob = get("ob_dim", None)
        if ob is None:
            ob = get("ob_dims")
        if isinstance(ob, (tuple, list)):
            ob = int(ob[-1])
        q_agg = get("q_agg", "mean")
        if q_agg not in ("mean", "min"):
            raise ValueError(f"unknown q_agg {q_agg!r}")
        self.P = len(alphas)
        assert len(seeds) == self.P
        c = Config(
            ob_dim=int(ob), action_dim=int(get("action_dim")), hidden_actor=ha, hidden_value=hv,
            n_hidden_actor=nha, n_hidden_value=nhv, num_ensembles=int(get("num_ensembles", 2)),
            flow_steps=int(get("flow_steps", 10)), batch_size=int(get("batch_size", 256)), population=self.P,
            lr=float(get("lr", 3e-4)), discount=float(get("discount", 0.99)), tau=float(get("tau", 0.005)),
            adam_b1=float(get("adam_b1", 0.9)), adam_b2=float(get("adam_b2", 0.999)),
            adam_eps=float(get("adam_eps", 1e-8)), ln_eps=float(get("ln_eps", 1e-6)),
            q_agg_min=int(q_agg == "min"), normalize_q_loss=int(bool(get("normalize_q_loss", False))),
            layer_norm=int(bool(get("layer_norm", True))), actor_layer_norm=int(bool(get("actor_layer_norm", False))),
        )
        self.cfg = c
        self.O, self.A, self.B, self.E = c.ob_dim, c.action_dim, c.batch_size, c.num_ensembles
        self.alphas = np.asarray(alphas, dtype=np.float32)
        self.seeds = np.asarray(seeds, dtype=np.uint64)
        h = _vp()
        check(lib.fqlpop_create(ctypes.byref(c), self.alphas.ctypes.data, self.seeds.ctypes.data, ctypes.byref(h)),
              "fqlpop_create")
        self._h = h
        self._keep = {}
        self.active = list(range(self.P))

    def close(self):
        if getattr(self, "_h", None):
            lib.fqlpop_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- data -------------------------------------------------------------------------------------------
    def set_dataset(self, observations, actions, next_observations, rewards, masks):
        """Binds device-resident fp32 arrays (rows = transitions)."""
        ts = []
        for x in (observations, actions, next_observations, rewards, masks):
            t = torch.as_tensor(x, dtype=torch.float32).to(self.device).contiguous()
            ts.append(t)
        n = ts[0].shape[0]
        assert ts[0].shape == (n, self.O) and ts[1].shape == (n, self.A) and ts[2].shape == (n, self.O)
        assert ts[3].numel() == n and ts[4].numel() == n
        self._keep["ds"] = ts
        check(lib.fqlpop_set_dataset(self._h, *[t.data_ptr() for t in ts], n), "fqlpop_set_dataset")

    def set_noise(self, indices=None, z_critic=None, x0=None, t=None, z_actor=None, z_mse=None):
        """Parity mode: explicit [P,B] row indices and [P,B,A] noise ([P,B] for t). None -> on-device Philox."""
        def dev(x, dt):
            return None if x is None else torch.as_tensor(x).to(self.device, dt).contiguous()
        ts = [dev(indices, torch.int64)] + [dev(x, torch.float32) for x in (z_critic, x0, t, z_actor, z_mse)]
        self._keep["noise"] = ts
        check(lib.fqlpop_set_noise(self._h, *[_ptr(x) for x in ts]), "fqlpop_set_noise")

    def set_active(self, members: Sequence[int]):
        m = np.asarray(list(members), dtype=np.int32)
        if [int(x) for x in m] == list(self.active):
            return  # unchanged: keep the captured step graphs (set_active invalidates and re-captures them)
        check(lib.fqlpop_set_active(self._h, m.ctypes.data, len(m)), "fqlpop_set_active")
        self.active = [int(x) for x in m]

    # ---- compute ----------------------------------------------------------------------------------------
    def step(self, n_steps: int = 1, stream: Optional[torch.cuda.Stream] = None):
        s = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        check(lib.fqlpop_step(self._h, int(n_steps), s), "fqlpop_step")

    def total_loss(self, stream: Optional[torch.cuda.Stream] = None):
        s = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        check(lib.fqlpop_total_loss(self._h, s), "fqlpop_total_loss")

    def info(self, with_grad_stats: bool = False) -> np.ndarray:
        out = np.zeros((self.P, INFO_SLOTS), dtype=np.float32)
        s = torch.cuda.current_stream(self.device).cuda_stream
        check(lib.fqlpop_info(self._h, out.ctypes.data, int(with_grad_stats), s), "fqlpop_info")
        return out

    def info_async(self, out_pinned: torch.Tensor, stream: Optional[torch.cuda.Stream] = None):
        """Enqueue the copy of the info block into `out_pinned` ([P, INFO_SLOTS] float32, pinned) without waiting; record
        an event on the stream afterwards and synchronise on it before reading."""
        assert out_pinned.is_pinned() and out_pinned.dtype == torch.float32 and out_pinned.numel() == self.P * INFO_SLOTS
        s = (stream or torch.cuda.current_stream(self.device)).cuda_stream
        check(lib.fqlpop_info_async(self._h, out_pinned.data_ptr(), s), "fqlpop_info_async")

    def info_dict(self, member: int, with_grad_stats: bool = False) -> Dict[str, float]:
        row = self.info(with_grad_stats)[member]
        n = len(INFO_KEYS) if with_grad_stats else 10
        return {k: float(row[i]) for i, k in enumerate(INFO_KEYS[:n])}

    def last_batch(self):
        """(rewards, masks) [P,B] of the batch the last step gathered (test hook for the on-device gather)."""
        r = np.zeros((self.P, self.B), np.float32)
        m = np.zeros((self.P, self.B), np.float32)
        check(lib.fqlpop_last_batch(self._h, r.ctypes.data, m.ctypes.data), "fqlpop_last_batch")
        return r, m

    def rollout(self, envmodel, init_obs, max_episode_steps, **kw):
        from . import envmodel as _em
        return _em.rollout(self, envmodel, init_obs, max_episode_steps, **kw)

    def sample_actions(self, member, observations, noise=None, seed=0):
        from . import envmodel as _em
        return _em.sample_actions(self, member, observations, noise, seed)

    @property
    def launches_per_step(self) -> int:
        return int(lib.fqlpop_launches_per_step(self._h))

    # ---- parameters ---------------------------------------------------------------------------------------
    def _net_shape(self, net: str):
        c = self.cfg
        if net == "critic":
            return c.num_ensembles, c.ob_dim + c.action_dim, c.hidden_value, c.n_hidden_value, 1, bool(c.layer_norm)
        if net == "actor_bc_flow":
            return 0, c.ob_dim + c.action_dim + 1, c.hidden_actor, c.n_hidden_actor, c.action_dim, bool(c.actor_layer_norm)
        if net == "actor_onestep_flow":
            return 0, c.ob_dim + c.action_dim, c.hidden_actor, c.n_hidden_actor, c.action_dim, bool(c.actor_layer_norm)
        raise KeyError(net)

    def _leaf(self, net, layer, kind, which, member, arr: np.ndarray, to_host: bool):
        check(lib.fqlpop_leaf(self._h, NETS[net], layer, kind, WHICH[which], member, arr.ctypes.data, int(to_host)),
              f"fqlpop_leaf({net},{layer},{kind},{which})")

    def get_tree(self, member: int, net: str, which: str = "params") -> Dict[str, Dict[str, np.ndarray]]:
        """One network of one member as a flax-style tree: Dense_i/{kernel,bias}, LayerNorm_i/{scale,bias}."""
        ens, din, H, nh, dout, ln = self._net_shape(net)
        lead = (ens,) if ens else ()
        nb = max(ens, 1)
        tree = {}
        for l in range(nh + 1):
            k_in = din if l == 0 else H
            k_out = dout if l == nh else H
            ker = np.zeros((nb, k_in, k_out), np.float32)
            bias = np.zeros((nb, k_out), np.float32)
            self._leaf(net, l, 0, which, member, ker, True)
            self._leaf(net, l, 1, which, member, bias, True)
            tree[f"Dense_{l}"] = {"kernel": ker.reshape(*lead, k_in, k_out), "bias": bias.reshape(*lead, k_out)}
            if ln and l < nh:
                sc = np.zeros((nb, H), np.float32)
                bi = np.zeros((nb, H), np.float32)
                self._leaf(net, l, 2, which, member, sc, True)
                self._leaf(net, l, 3, which, member, bi, True)
                tree[f"LayerNorm_{l}"] = {"scale": sc.reshape(*lead, H), "bias": bi.reshape(*lead, H)}
        return tree

    def set_tree(self, member: int, net: str, tree, which: str = "params"):
        ens, din, H, nh, dout, ln = self._net_shape(net)
        for l in range(nh + 1):
            d = tree[f"Dense_{l}"]
            ker = np.ascontiguousarray(np.asarray(d["kernel"], dtype=np.float32))
            bias = np.ascontiguousarray(np.asarray(d["bias"], dtype=np.float32))
            k_in = din if l == 0 else H
            k_out = dout if l == nh else H
            assert ker.size == max(ens, 1) * k_in * k_out, (net, l, ker.shape)
            self._leaf(net, l, 0, which, member, ker, False)
            self._leaf(net, l, 1, which, member, bias, False)
            if ln and l < nh:
                n = tree[f"LayerNorm_{l}"]
                self._leaf(net, l, 2, which, member, np.ascontiguousarray(np.asarray(n["scale"], np.float32)), False)
                self._leaf(net, l, 3, which, member, np.ascontiguousarray(np.asarray(n["bias"], np.float32)), False)

    def refresh_shadow(self):
        check(lib.fqlpop_refresh_shadow(self._h, torch.cuda.current_stream(self.device).cuda_stream),
              "fqlpop_refresh_shadow")
        torch.cuda.synchronize(self.device)

    def get_counters(self):
        """(adam_steps [P] int32, step_counter [P] uint64): optimiser step counts and each member's Philox stream position."""
        steps = np.zeros(self.P, np.int32)
        ctr = np.zeros(self.P, np.uint64)
        check(lib.fqlpop_counters(self._h, steps.ctypes.data, ctr.ctypes.data, 0), "fqlpop_counters")
        return steps, ctr

    def set_counters(self, adam_steps=None, step_counter=None):
        steps = None if adam_steps is None else np.ascontiguousarray(np.asarray(adam_steps, np.int32))
        ctr = None if step_counter is None else np.ascontiguousarray(
            np.broadcast_to(np.asarray(step_counter, np.uint64), (self.P,)))
        check(lib.fqlpop_counters(self._h, None if steps is None else steps.ctypes.data,
                                  None if ctr is None else ctr.ctypes.data, 1), "fqlpop_counters")


    # ---- survivor migration (include/fqlpop.h: fqlpop_member_export / _import) -----------------------------------
    def member_state_words(self) -> int:
        return int(lib.fqlpop_member_state_words(self._h))

    def export_member(self, member: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Slot `member`'s whole training state as one float32 device tensor (a bit container: the header words are
        integers), ready for torch.distributed.send / recv. Stream-ordered on the current stream."""
        n = self.member_state_words()
        if out is None:
            out = torch.empty(n, dtype=torch.float32, device=self.device)
        assert out.is_cuda and out.dtype == torch.float32 and out.numel() == n and out.is_contiguous()
        check(lib.fqlpop_member_export(self._h, int(member), out.data_ptr(),
                                       torch.cuda.current_stream(self.device).cuda_stream), "fqlpop_member_export")
        return out

    def import_member(self, member: int, blob: torch.Tensor):
        """Overwrites slot `member` with a state exported by a population of the same network shapes (alpha, seed, Adam
        step and Philox position travel with it)."""
        assert blob.is_cuda and blob.dtype == torch.float32 and blob.is_contiguous()
        assert blob.numel() == self.member_state_words(), "blob of another population shape"
        check(lib.fqlpop_member_import(self._h, int(member), blob.data_ptr(),
                                       torch.cuda.current_stream(self.device).cuda_stream), "fqlpop_member_import")
        head = blob[:4].view(torch.int32).cpu().numpy().view(np.uint32)
        self.alphas[member] = head[1:2].view(np.float32)[0]
        self.seeds[member] = np.uint64(head[2]) | (np.uint64(head[3]) << np.uint64(32))


def philox_fill(seed: int, step: int, rows: int, slots: int, kind: int, n: int = 0) -> np.ndarray:
    """Device Philox stream (see include/fqlpop.h fqlpop_philox_fill)."""
    if kind == 0:
        out = np.zeros((rows, slots, 4), np.uint32)
    elif kind == 2:
        out = np.zeros((rows, slots), np.uint32)
    else:
        out = np.zeros((rows, slots), np.float32)
    check(lib.fqlpop_philox_fill(seed, step, rows, slots, kind, n, out.ctypes.data), "fqlpop_philox_fill")
    return out
