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
