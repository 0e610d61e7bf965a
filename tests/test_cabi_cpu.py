"""The C-ABI library loads without a GPU and exports every symbol include/fqlpop.h declares; entry points fail
loudly (status != 0 + message) instead of falling back to the CPU."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "fqlpop.h")
LIB = os.path.join(ROOT, "flow-q-learning_b200", "lib", "libfqlpop.so")


def declared_symbols():
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    return sorted(set(re.findall(r"\b(fqlpop_[a-z_0-9]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    names = declared_symbols()
    for must in ("fqlpop_create", "fqlpop_step", "fqlpop_total_loss", "fqlpop_info", "fqlpop_set_dataset",
                 "fqlpop_rollout", "fqlpop_sample_actions", "fqlpop_env_step", "fqlpop_envmodel_create", "fqlpop_leaf"):
        assert must in names
    assert len(names) >= 20


def test_library_exports_every_declared_symbol():
    assert os.path.exists(LIB), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(LIB)
    missing = [n for n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, missing
    lib.fqlpop_version.restype = ctypes.c_int
    assert lib.fqlpop_version() >= 1


def test_no_silent_cpu_fallback_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("GPU present")
    from fqlpop import _lib
    from fqlpop.population import Population
    import pytest
    with pytest.raises(_lib.FqlPopError):
        Population(dict(ob_dim=28, action_dim=5), [10.0], [0])


def test_argument_errors_come_back_as_status_and_message():
    """Error behaviour of the boundary: a non-zero status and a message from fqlpop_last_error, never an exception or
    a crash; checked on the host-only argument validation (no device work)."""
    lib = ctypes.CDLL(LIB)
    lib.fqlpop_last_error.restype = ctypes.c_char_p
    vp = ctypes.c_void_p
    lib.fqlpop_step.argtypes = [vp, ctypes.c_int32, vp]
    assert lib.fqlpop_step(None, 1, None) != 0 and lib.fqlpop_last_error()
    lib.fqlpop_wm_step.argtypes = [vp, ctypes.c_int32] + [vp] * 6
    assert lib.fqlpop_wm_step(None, 1, None, None, None, None, None, None) != 0
    assert b"handle" in lib.fqlpop_last_error()
    lib.fqlpop_decoder_sample.argtypes = [vp, ctypes.c_int32, vp, ctypes.c_uint64, vp, vp]
    assert lib.fqlpop_decoder_sample(None, 4, None, 0, None, None) != 0
    hid = (ctypes.c_int32 * 1)(64)
    out = vp()
    lib.fqlpop_decoder_create.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.POINTER(ctypes.c_int32), ctypes.c_int32,
                                          vp, ctypes.POINTER(vp)]
    assert lib.fqlpop_decoder_create(8, 0, hid, 28, None, ctypes.byref(out)) != 0   # no hidden layer
    assert b"hidden" in lib.fqlpop_last_error()
    from fqlpop.worldmodel import WmConfig
    lib.fqlpop_wm_num_params.argtypes = [ctypes.POINTER(WmConfig)]
    lib.fqlpop_wm_num_params.restype = ctypes.c_int64
    c = WmConfig()
    c.n_hidden = 0
    assert lib.fqlpop_wm_num_params(ctypes.byref(c)) == -1
    # VAE trainer: null handles, a configuration without hidden layers
    from fqlpop.worldmodel import VaeConfig
    lib.fqlpop_vae_step.argtypes = [vp, ctypes.c_int32, vp, vp, vp]
    assert lib.fqlpop_vae_step(None, 1, None, None, None) != 0 and b"handle" in lib.fqlpop_last_error()
    lib.fqlpop_vae_train.argtypes = [vp, ctypes.c_int32, vp]
    assert lib.fqlpop_vae_train(None, 1, None) != 0
    lib.fqlpop_vae_num_params.argtypes = [ctypes.POINTER(VaeConfig)]
    lib.fqlpop_vae_num_params.restype = ctypes.c_int64
    v = VaeConfig()
    v.ob_dim, v.latent_dim, v.batch, v.n_hidden = 28, 4, 64, 0
    assert lib.fqlpop_vae_num_params(ctypes.byref(v)) == -1
    lib.fqlpop_vae_create.argtypes = [ctypes.POINTER(VaeConfig), vp, ctypes.POINTER(vp)]
    assert lib.fqlpop_vae_create(ctypes.byref(v), None, ctypes.byref(out)) != 0 and b"hidden" in lib.fqlpop_last_error()
    c.kind, c.ob_dim, c.action_dim, c.n_hidden = 2, 42, 8, 3
    for i, h in enumerate((128, 256, 128)):
        c.hidden[i] = h
    # multistep shares the state predictor's layout: LayerNorm (2 * 50) + 50*128+128 + 128*256+256 + 256*128+128 + 128*42+42
    assert lib.fqlpop_wm_num_params(ctypes.byref(c)) == 100 + 6528 + 33024 + 32896 + 5418
    lib.fqlpop_wm_create.argtypes = [ctypes.POINTER(WmConfig), vp, vp, ctypes.POINTER(vp)]
    assert lib.fqlpop_wm_create(ctypes.byref(c), None, None, ctypes.byref(out)) != 0   # decay_steps 0 / no parameters
