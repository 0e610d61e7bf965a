"""ctypes binding of the world-model training entry points of libfqlpop.so (include/fqlpop.h, fqlpop_wm_*).

Replaces the jitted train_step / eval_step of StatePredictorTrainer
(/root/reference/envmodel/state_predictor_trainer.py:70-125) and TerminationPredictorTrainer
(/root/reference/envmodel/termination_predictor_trainer.py:58-126). Parameter trees use flax's auto-names
(`LayerNorm_0`, `Dense_i` with `kernel` [in][out] / `bias`), i.e. what `flax.serialization.to_bytes(state.params)`
stores under "params" (train_env_model.py:123-127) and what utils/envmodel.load_model reads back.
There is no CPU fallback.
"""
import ctypes
from typing import Dict, Optional, Sequence

import numpy as np
import torch

from ._lib import check, default_device, lib

STATE, TERMINATION, MULTISTEP = 0, 1, 2
LOG_SLOTS = 16
_STATE_LOGS = {"loss": 0, "next_observation_loss": 1, "termination_loss": 2, "true_termination_loss": 3,
               "false_termination_loss": 4}
_TERM_LOGS = {"loss": 0, "true_loss": 3, "false_loss": 4}
_TERM_EVAL_LOGS = {"accuracy": 6, "precision": 7, "recall": 8, "f1": 9}


class WmConfig(ctypes.Structure):
    _fields_ = [
        ("kind", ctypes.c_int32), ("ob_dim", ctypes.c_int32), ("action_dim", ctypes.c_int32), ("batch", ctypes.c_int32),
        ("n_hidden", ctypes.c_int32), ("hidden", ctypes.c_int32 * 8),
        ("term_n_hidden", ctypes.c_int32), ("term_hidden", ctypes.c_int32 * 8),
        ("decay_steps", ctypes.c_int32), ("lr", ctypes.c_float), ("ln_eps", ctypes.c_float),
        ("true_termination_weight", ctypes.c_float), ("termination_weight", ctypes.c_float),
        ("reconstruction_weight", ctypes.c_float),
        ("focal_alpha", ctypes.c_float), ("focal_gamma", ctypes.c_float),
        ("seed", ctypes.c_uint64), ("seq_len", ctypes.c_int32), ("reserved", ctypes.c_int32),
    ]


_F = ctypes.POINTER(ctypes.c_float)
lib.fqlpop_wm_num_params.argtypes = [ctypes.POINTER(WmConfig)]
lib.fqlpop_wm_num_params.restype = ctypes.c_int64
lib.fqlpop_wm_create.argtypes = [ctypes.POINTER(WmConfig), ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
lib.fqlpop_wm_destroy.argtypes = [ctypes.c_void_p]
lib.fqlpop_wm_step.argtypes = [ctypes.c_void_p, ctypes.c_int32] + [ctypes.c_void_p] * 6
lib.fqlpop_wm_set_dataset.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int64]
lib.fqlpop_wm_train.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p]
lib.fqlpop_wm_logs.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
lib.fqlpop_wm_export.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
lib.fqlpop_wm_launches.argtypes = [ctypes.c_void_p]
lib.fqlpop_wm_launches.restype = ctypes.c_int64


def _np32(x) -> np.ndarray:
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.ascontiguousarray(np.asarray(x, dtype=np.float32))


def layer_dims(kind: int, ob_dim: int, action_dim: int, hidden: Sequence[int]):
    return ([ob_dim + action_dim, *hidden, ob_dim] if kind != TERMINATION else [ob_dim, *hidden, 1])


def flatten_params(tree: Dict, kind: int) -> np.ndarray:
    """flax tree -> the flat vector of include/fqlpop.h (LayerNorm scale, bias, then kernel, bias per Dense layer)."""
    parts = []
    if kind != TERMINATION:
        parts += [_np32(tree["LayerNorm_0"]["scale"]).ravel(), _np32(tree["LayerNorm_0"]["bias"]).ravel()]
    n = sum(1 for k in tree if k.startswith("Dense_"))
    for i in range(n):
        parts += [_np32(tree[f"Dense_{i}"]["kernel"]).ravel(), _np32(tree[f"Dense_{i}"]["bias"]).ravel()]
    return np.ascontiguousarray(np.concatenate(parts))


def unflatten_params(flat: np.ndarray, kind: int, dims: Sequence[int]) -> Dict:
    tree, off = {}, 0

    def take(*shape):
        nonlocal off
        n = int(np.prod(shape))
        out = flat[off:off + n].reshape(shape).copy()
        off += n
        return out

    if kind != TERMINATION:
        tree["LayerNorm_0"] = {"scale": take(dims[0]), "bias": take(dims[0])}
    for i in range(len(dims) - 1):
        tree[f"Dense_{i}"] = {"kernel": take(dims[i], dims[i + 1]), "bias": take(dims[i + 1])}
    assert off == flat.size
    return tree


def init_params(seed: int, kind: int, ob_dim: int, action_dim: int, hidden: Sequence[int]) -> Dict:
    """flax defaults for nn.Dense / nn.LayerNorm without initialisers (envmodel/baseline.py:31-36): LeCun-normal kernels
    (truncated at two standard deviations, variance 1/fan_in), zero biases, scale 1. The random stream is numpy's, not
    JAX's threefry."""
    rng = np.random.default_rng(seed)
    dims = layer_dims(kind, ob_dim, action_dim, hidden)
    tree = {}
    if kind != TERMINATION:
        tree["LayerNorm_0"] = {"scale": np.ones(dims[0], np.float32), "bias": np.zeros(dims[0], np.float32)}
    for i in range(len(dims) - 1):
        n = dims[i] * dims[i + 1]
        z = rng.standard_normal(n)
        while True:  # truncated normal on [-2, 2] by resampling
            bad = np.abs(z) > 2.0
            if not bad.any():
                break
            z[bad] = rng.standard_normal(int(bad.sum()))
        std = np.sqrt(1.0 / dims[i]) / 0.87962566103423978
        tree[f"Dense_{i}"] = {"kernel": (z * std).astype(np.float32).reshape(dims[i], dims[i + 1]),
                              "bias": np.zeros(dims[i + 1], np.float32)}
    return tree


class WorldModelTrainer:
    """Device-resident trainer of one world-model network. `step` takes an explicit batch (parity mode, validation);
    `set_dataset` + `train(n)` keep the dataset in HBM and sample on the device."""

    def __init__(self, kind: int, ob_dim: int, action_dim: int, hidden_dims: Sequence[int], batch_size: int, steps: int,
                 init_learning_rate: float = 1e-3, params: Optional[Dict] = None, seed: int = 0,
                 termination_params: Optional[Dict] = None, termination_hidden_dims: Sequence[int] = (),
                 true_termination_weight: float = 30.0, termination_weight: float = 1.0,
                 reconstruction_weight: float = 0.0, focal_alpha: float = 0.25, focal_gamma: float = 2.0,
                 ln_eps: float = 1e-6, sequence_length: int = 1, device=None):
        self.device = torch.device(device) if device is not None else default_device()
        self.kind, self.ob_dim, self.action_dim = kind, ob_dim, action_dim
        self.sequence_length = int(sequence_length) if kind == MULTISTEP else 1
        self.hidden_dims, self.batch_size = tuple(int(h) for h in hidden_dims), int(batch_size)
        self.dims = layer_dims(kind, ob_dim, action_dim, self.hidden_dims)
        c = WmConfig()
        c.kind, c.ob_dim, c.action_dim, c.batch = kind, ob_dim, action_dim, batch_size
        c.n_hidden = len(self.hidden_dims)
        for i, h in enumerate(self.hidden_dims):
            c.hidden[i] = h
        with_term = kind != TERMINATION and termination_weight > 0
        if with_term:
            if termination_params is None:
                raise ValueError("termination_weight > 0 needs the trained termination predictor's parameters")
            c.term_n_hidden = len(termination_hidden_dims)
            for i, h in enumerate(termination_hidden_dims):
                c.term_hidden[i] = int(h)
        c.decay_steps, c.lr, c.ln_eps = int(steps), init_learning_rate, ln_eps
        c.true_termination_weight, c.termination_weight = true_termination_weight, termination_weight
        c.reconstruction_weight = reconstruction_weight
        c.focal_alpha, c.focal_gamma, c.seed = focal_alpha, focal_gamma, seed
        c.seq_len = self.sequence_length
        self._cfg = c
        self.n_params = int(lib.fqlpop_wm_num_params(ctypes.byref(c)))
        if params is None:
            params = init_params(seed, kind, ob_dim, action_dim, self.hidden_dims)
        flat = flatten_params(params, kind)
        if flat.size != self.n_params:
            raise ValueError(f"parameter tree has {flat.size} values, the network needs {self.n_params}")
        tflat = flatten_params(termination_params, TERMINATION) if with_term else None
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(lib.fqlpop_wm_create(ctypes.byref(c), flat.ctypes.data, tflat.ctypes.data if with_term else None,
                                       ctypes.byref(self._h)), "fqlpop_wm_create")
        self._keep = {}

    def close(self):
        if getattr(self, "_h", None):
            lib.fqlpop_wm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _dev(self, x):
        return torch.as_tensor(np.array(x, dtype=np.float32) if not torch.is_tensor(x) else x,
                               dtype=torch.float32).to(self.device).contiguous()

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def logs(self) -> np.ndarray:
        out = np.zeros(LOG_SLOTS, np.float32)
        check(lib.fqlpop_wm_logs(self._h, out.ctypes.data, self._stream()), "fqlpop_wm_logs")
        return out

    def named_logs(self, train: bool) -> Dict[str, float]:
        raw = self.logs()
        if self.kind != TERMINATION:
            keys = dict(_STATE_LOGS)
            if not self._cfg.termination_weight > 0:
                keys = {"loss": 0, "next_observation_loss": 1}
            if self._cfg.reconstruction_weight > 0:
                keys["reconstruction_loss"] = 5
        else:
            keys = dict(_TERM_LOGS)
            if not train:
                keys.update(_TERM_EVAL_LOGS)
        return {k: float(raw[i]) for k, i in keys.items()}

    def step(self, batch: Dict, train: bool = True, dropout_mask=None) -> Dict[str, float]:
        """One train_step (train=True) or eval_step on an explicit batch; returns the reference's logs dict."""
        with torch.cuda.device(self.device):
            t = {k: self._dev(batch[k]) for k in ("observations", "actions", "next_observations", "rewards") if k in batch}
            n = t["next_observations"].shape[0]
            if n != self.batch_size:
                raise ValueError(f"batch has {n} rows, the trainer was built for {self.batch_size}")
            if self.kind == MULTISTEP and tuple(t["actions"].shape[:2]) != (n, self.sequence_length):
                raise ValueError(f"multistep windows must be [batch, {self.sequence_length}, ...]")
            m = self._dev(dropout_mask) if (dropout_mask is not None and train) else None
            ptr = lambda k: t[k].data_ptr() if k in t else None
            check(lib.fqlpop_wm_step(self._h, 1 if train else 0, ptr("observations"), ptr("actions"),
                                     ptr("next_observations"), ptr("rewards"), m.data_ptr() if m is not None else None,
                                     self._stream()), "fqlpop_wm_step")
            return self.named_logs(train)

    def set_dataset(self, dataset: Dict):
        with torch.cuda.device(self.device):
            self._keep = {k: self._dev(dataset[k]) for k in ("observations", "actions", "next_observations", "rewards")}
            d = self._keep
            check(lib.fqlpop_wm_set_dataset(self._h, d["observations"].data_ptr(), d["actions"].data_ptr(),
                                            d["next_observations"].data_ptr(), d["rewards"].data_ptr(),
                                            d["observations"].shape[0]), "fqlpop_wm_set_dataset")

    def train(self, n_steps: int, dropout_mask=None):
        """n_steps optimiser steps on batches sampled on the device from the resident dataset (asynchronous)."""
        with torch.cuda.device(self.device):
            if dropout_mask is not None:
                self._keep["mask"] = self._dev(dropout_mask)
            m = self._keep.get("mask")
            check(lib.fqlpop_wm_train(self._h, int(n_steps), m.data_ptr() if m is not None else None, self._stream()),
                  "fqlpop_wm_train")

    def export(self, which: int = 0) -> np.ndarray:
        out = np.zeros(self.n_params, np.float32)
        check(lib.fqlpop_wm_export(self._h, which, out.ctypes.data, self.n_params, self._stream()), "fqlpop_wm_export")
        return out

    def params(self) -> Dict:
        return unflatten_params(self.export(0), self.kind, self.dims)

    def grads(self) -> Dict:
        return unflatten_params(self.export(3), self.kind, self.dims)

    @property
    def launches(self) -> int:
        return int(lib.fqlpop_wm_launches(self._h))


lib.fqlpop_decoder_create.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.POINTER(ctypes.c_int32), ctypes.c_int32,
                                      ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
lib.fqlpop_decoder_destroy.argtypes = [ctypes.c_void_p]
lib.fqlpop_decoder_sample.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p,
                                      ctypes.c_void_p]


class Decoder:
    """Device copy of the VAE decoder that generates initial observations (envmodel/initial_observation.py:21-30):
    `sample(n, seed)` = decoder(z), z ~ N(0, I) -- the sampling of distribution_visualizer.py:96-99, on the GPU, so that a
    simulated evaluation never touches a real environment. `params` is the flax subtree ["params"]["decoder"]."""

    def __init__(self, params: Dict, latent_dim: int, hidden_dims: Sequence[int], output_dimension: int, device=None):
        self.device = torch.device(device) if device is not None else default_device()
        self.latent_dim, self.output_dimension = int(latent_dim), int(output_dimension)
        self.hidden_dims = tuple(int(h) for h in hidden_dims)
        flat = flatten_params(params, TERMINATION)  # plain Dense stack: kernel, bias per layer
        dims = [self.latent_dim, *self.hidden_dims, self.output_dimension]
        want = sum(a * b + b for a, b in zip(dims[:-1], dims[1:]))
        if flat.size != want:
            raise ValueError(f"decoder parameters hold {flat.size} values, dims {dims} need {want}")
        hid = (ctypes.c_int32 * len(self.hidden_dims))(*self.hidden_dims)
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(lib.fqlpop_decoder_create(self.latent_dim, len(self.hidden_dims), hid, self.output_dimension,
                                            flat.ctypes.data, ctypes.byref(self._h)), "fqlpop_decoder_create")

    def sample(self, n: int, seed: int = 0, noise=None):
        """-> torch.float32 [n, output_dimension] on the device."""
        with torch.cuda.device(self.device):
            out = torch.empty(n, self.output_dimension, dtype=torch.float32, device=self.device)
            z = None
            if noise is not None:
                z = torch.as_tensor(noise, dtype=torch.float32).to(self.device).contiguous()
                if tuple(z.shape) != (n, self.latent_dim):
                    raise ValueError(f"noise must be [{n}, {self.latent_dim}]")
            check(lib.fqlpop_decoder_sample(self._h, n, z.data_ptr() if z is not None else None, int(seed), out.data_ptr(),
                                            torch.cuda.current_stream(self.device).cuda_stream), "fqlpop_decoder_sample")
            return out

    def close(self):
        if getattr(self, "_h", None):
            lib.fqlpop_decoder_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- initial-observation VAE training (include/fqlpop.h: fqlpop_vae_*) ------------------------------------------------
class VaeConfig(ctypes.Structure):
    _fields_ = [("ob_dim", ctypes.c_int32), ("latent_dim", ctypes.c_int32), ("batch", ctypes.c_int32),
                ("n_hidden", ctypes.c_int32), ("hidden", ctypes.c_int32 * 8), ("decay_steps", ctypes.c_int32),
                ("lr", ctypes.c_float), ("reconstruction_weight", ctypes.c_float), ("seed", ctypes.c_uint64)]


lib.fqlpop_vae_num_params.argtypes = [ctypes.POINTER(VaeConfig)]
lib.fqlpop_vae_num_params.restype = ctypes.c_int64
lib.fqlpop_vae_create.argtypes = [ctypes.POINTER(VaeConfig), ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]
lib.fqlpop_vae_destroy.argtypes = [ctypes.c_void_p]
lib.fqlpop_vae_step.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
lib.fqlpop_vae_set_dataset.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64]
lib.fqlpop_vae_train.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p]
lib.fqlpop_vae_logs.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
lib.fqlpop_vae_export.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
lib.fqlpop_vae_launches.argtypes = [ctypes.c_void_p]
lib.fqlpop_vae_launches.restype = ctypes.c_int64

_VAE_LOGS = {"reconstruction_loss": 5, "kl": 12, "loss": 0}   # the keys of vae_loss's logs (initial_observation.py:71-75)


def vae_dims(ob_dim: int, latent_dim: int, hidden: Sequence[int]):
    hidden = [int(h) for h in hidden]
    return [ob_dim, *hidden, 2 * latent_dim], [latent_dim, *hidden[::-1], ob_dim]


def _dense_tree(rng, dims):
    tree = {}
    for i in range(len(dims) - 1):
        z = rng.standard_normal(dims[i] * dims[i + 1])
        while True:  # LeCun normal truncated at two standard deviations (flax's nn.Dense default), by resampling
            bad = np.abs(z) > 2.0
            if not bad.any():
                break
            z[bad] = rng.standard_normal(int(bad.sum()))
        std = np.sqrt(1.0 / dims[i]) / 0.87962566103423978
        tree[f"Dense_{i}"] = {"kernel": (z * std).astype(np.float32).reshape(dims[i], dims[i + 1]),
                              "bias": np.zeros(dims[i + 1], np.float32)}
    return tree


def init_vae_params(seed: int, ob_dim: int, latent_dim: int, hidden: Sequence[int]) -> Dict:
    """{"encoder": Dense_0 .. Dense_{h-1}, Dense_h (mean), Dense_{h+1} (logvar); "decoder": Dense_0 .. Dense_h} with flax's
    defaults for nn.Dense (envmodel/initial_observation.py:15-18, 27-29). numpy's random stream, not JAX's."""
    rng = np.random.default_rng(seed)
    enc_dims, dec_dims = vae_dims(ob_dim, latent_dim, hidden)
    h = len(hidden)
    enc = _dense_tree(rng, enc_dims[:-1] + [latent_dim])                        # hidden layers + the mean head
    enc[f"Dense_{h + 1}"] = _dense_tree(rng, [enc_dims[-2], latent_dim])["Dense_0"]   # the logvar head
    return {"encoder": enc, "decoder": _dense_tree(rng, dec_dims)}


def flatten_vae_params(tree: Dict, n_hidden: int) -> np.ndarray:
    """flax tree -> the flat vector of include/fqlpop.h: the mean and logvar heads become ONE layer of 2 L columns."""
    enc, dec, parts = tree["encoder"], tree["decoder"], []
    for i in range(n_hidden):
        parts += [_np32(enc[f"Dense_{i}"]["kernel"]).ravel(), _np32(enc[f"Dense_{i}"]["bias"]).ravel()]
    mean, logvar = enc[f"Dense_{n_hidden}"], enc[f"Dense_{n_hidden + 1}"]
    parts += [np.concatenate([_np32(mean["kernel"]), _np32(logvar["kernel"])], axis=1).ravel(),
              np.concatenate([_np32(mean["bias"]), _np32(logvar["bias"])])]
    for i in range(n_hidden + 1):
        parts += [_np32(dec[f"Dense_{i}"]["kernel"]).ravel(), _np32(dec[f"Dense_{i}"]["bias"]).ravel()]
    return np.ascontiguousarray(np.concatenate(parts))


def unflatten_vae_params(flat: np.ndarray, ob_dim: int, latent_dim: int, hidden: Sequence[int]) -> Dict:
    enc_dims, dec_dims = vae_dims(ob_dim, latent_dim, hidden)
    off = 0

    def take(*shape):
        nonlocal off
        n = int(np.prod(shape))
        out = flat[off:off + n].reshape(shape).copy()
        off += n
        return out

    h, L = len(hidden), latent_dim
    enc = {}
    for i in range(h):
        enc[f"Dense_{i}"] = {"kernel": take(enc_dims[i], enc_dims[i + 1]), "bias": take(enc_dims[i + 1])}
    k, b = take(enc_dims[h], 2 * L), take(2 * L)
    enc[f"Dense_{h}"] = {"kernel": np.ascontiguousarray(k[:, :L]), "bias": b[:L].copy()}
    enc[f"Dense_{h + 1}"] = {"kernel": np.ascontiguousarray(k[:, L:]), "bias": b[L:].copy()}
    dec = {f"Dense_{i}": {"kernel": take(dec_dims[i], dec_dims[i + 1]), "bias": take(dec_dims[i + 1])}
           for i in range(h + 1)}
    assert off == flat.size
    return {"encoder": enc, "decoder": dec}


class VaeTrainer:
    """Device-resident trainer of the initial-observation VAE (envmodel/initial_observation.py:32-78). `step` takes an
    explicit batch (and, for parity tests, the reparameterisation noise); `set_dataset` + `train(n)` keep the observations
    in HBM and sample batches and noise on the device."""

    def __init__(self, ob_dim: int, latent_dim: int, hidden_dims: Sequence[int], batch_size: int, steps: int,
                 init_learning_rate: float = 1e-3, reconstruction_weight: float = 1.0, params: Optional[Dict] = None,
                 seed: int = 0, device=None):
        self.device = torch.device(device) if device is not None else default_device()
        self.ob_dim, self.latent_dim = int(ob_dim), int(latent_dim)
        self.hidden_dims, self.batch_size = tuple(int(h) for h in hidden_dims), int(batch_size)
        c = VaeConfig()
        c.ob_dim, c.latent_dim, c.batch, c.n_hidden = self.ob_dim, self.latent_dim, self.batch_size, len(self.hidden_dims)
        if not 1 <= c.n_hidden <= 7:
            raise ValueError("the VAE needs between 1 and 7 hidden layers")
        for i, h in enumerate(self.hidden_dims):
            c.hidden[i] = h
        c.decay_steps, c.lr, c.reconstruction_weight, c.seed = int(steps), init_learning_rate, reconstruction_weight, seed
        self._cfg = c
        self.n_params = int(lib.fqlpop_vae_num_params(ctypes.byref(c)))
        if params is None:
            params = init_vae_params(seed, self.ob_dim, self.latent_dim, self.hidden_dims)
        flat = flatten_vae_params(params, len(self.hidden_dims))
        if flat.size != self.n_params:
            raise ValueError(f"parameter tree has {flat.size} values, the VAE needs {self.n_params}")
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(lib.fqlpop_vae_create(ctypes.byref(c), flat.ctypes.data, ctypes.byref(self._h)), "fqlpop_vae_create")
        self._keep = {}

    def close(self):
        if getattr(self, "_h", None):
            lib.fqlpop_vae_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _dev(self, x):
        return torch.as_tensor(np.array(x, dtype=np.float32) if not torch.is_tensor(x) else x,
                               dtype=torch.float32).to(self.device).contiguous()

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def logs(self) -> np.ndarray:
        out = np.zeros(LOG_SLOTS, np.float32)
        check(lib.fqlpop_vae_logs(self._h, out.ctypes.data, self._stream()), "fqlpop_vae_logs")
        return out

    def named_logs(self) -> Dict[str, float]:
        raw = self.logs()
        return {k: float(raw[i]) for k, i in _VAE_LOGS.items()}

    def step(self, observations, eps=None, train: bool = True) -> Dict[str, float]:
        with torch.cuda.device(self.device):
            o = self._dev(observations)
            if tuple(o.shape) != (self.batch_size, self.ob_dim):
                raise ValueError(f"batch must be [{self.batch_size}, {self.ob_dim}], got {tuple(o.shape)}")
            e = self._dev(eps) if eps is not None else None
            if e is not None and tuple(e.shape) != (self.batch_size, self.latent_dim):
                raise ValueError(f"eps must be [{self.batch_size}, {self.latent_dim}]")
            self._keep["step"] = (o, e)   # the captured graph holds these pointers until the call has run
            check(lib.fqlpop_vae_step(self._h, 1 if train else 0, o.data_ptr(), e.data_ptr() if e is not None else None,
                                      self._stream()), "fqlpop_vae_step")
            return self.named_logs()

    def set_dataset(self, observations):
        with torch.cuda.device(self.device):
            self._keep["ds"] = self._dev(observations)
            check(lib.fqlpop_vae_set_dataset(self._h, self._keep["ds"].data_ptr(), self._keep["ds"].shape[0]),
                  "fqlpop_vae_set_dataset")

    def train(self, n_steps: int):
        """n_steps optimiser steps on batches (and latent noise) drawn on the device (asynchronous)."""
        with torch.cuda.device(self.device):
            check(lib.fqlpop_vae_train(self._h, int(n_steps), self._stream()), "fqlpop_vae_train")

    def export(self, which: int = 0) -> np.ndarray:
        out = np.zeros(self.n_params, np.float32)
        check(lib.fqlpop_vae_export(self._h, which, out.ctypes.data, self.n_params, self._stream()), "fqlpop_vae_export")
        return out

    def params(self) -> Dict:
        return unflatten_vae_params(self.export(0), self.ob_dim, self.latent_dim, self.hidden_dims)

    def grads(self) -> Dict:
        return unflatten_vae_params(self.export(3), self.ob_dim, self.latent_dim, self.hidden_dims)

    @property
    def launches(self) -> int:
        return int(lib.fqlpop_vae_launches(self._h))
