"""Survivor migration on the device (include/fqlpop.h: fqlpop_member_export / _import): a member exported from
one population and imported into a slot of another continues EXACTLY as it would have where it was -- same batches, same
noise (its Philox position travels with it), same parameters bit for bit -- and a blob of a population with other
network shapes is refused without touching anything."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import fql_oracle as orc  # noqa: E402  (only for the config dataclass)
from test_update_gpu import _cfg_dict  # noqa: E402

NETS = ("critic", "actor_bc_flow", "actor_onestep_flow")


def _dataset(cfg, n=4096, seed=3):
    g = torch.Generator().manual_seed(seed)
    O, A = cfg.ob_dim, cfg.action_dim
    obs = torch.randn(n, O, generator=g)
    return (obs, (torch.rand(n, A, generator=g) * 2 - 1) * 0.99, obs + 0.05 * torch.randn(n, O, generator=g),
            -(torch.rand(n, generator=g) < 0.9).float(), (torch.rand(n, generator=g) < 0.9).float())


def _state(pop, m):
    out = {}
    for net in NETS:
        for which in ("params", "mu", "nu") + (("target",) if net == "critic" else ()):
            for layer, leaves in pop.get_tree(m, net, which).items():
                for leaf, v in leaves.items():
                    out[f"{net}/{which}/{layer}/{leaf}"] = v
    return out


def test_migrated_member_continues_bit_for_bit():
    from fqlpop.population import Population
    cfg = orc.FQLConfig(ob_dim=28, action_dim=5, batch_size=64)
    data = _dataset(cfg)
    home = Population(_cfg_dict(cfg), [3.0, 30.0, 300.0], [11, 12, 13])
    away = Population(_cfg_dict(cfg), [1.0, 1.0], [0, 0])            # two placeholder slots, as the Trainer's spares
    for pop in (home, away):
        pop.set_dataset(*data)
        pop.set_noise()
    home.step(3)                                                     # the member has a history: Adam moments, counters
    blob = home.export_member(1)
    assert blob.numel() == home.member_state_words() == away.member_state_words()
    away.import_member(1, blob)
    assert float(away.alphas[1]) == 30.0 and int(away.seeds[1]) == 12
    before_h, before_a = _state(home, 1), _state(away, 1)
    for k in before_h:
        np.testing.assert_array_equal(before_h[k], before_a[k], err_msg=k)
    sh, ch = home.get_counters()
    sa, ca = away.get_counters()
    assert sa[1] == sh[1] == 3 and ca[1] == ch[1] and sa[0] == 0
    # both copies train on alone, each on "its" GPU slot: the same kernels see the same inputs
    home.set_active([1])
    away.set_active([1])
    home.step(4)
    away.step(4)
    torch.cuda.synchronize()
    after_h, after_a = _state(home, 1), _state(away, 1)
    changed = 0
    for k in after_h:
        np.testing.assert_array_equal(after_h[k], after_a[k], err_msg=k)
        changed += int(not np.array_equal(after_h[k], before_h[k]))
    assert changed > len(after_h) // 2                               # (they did train)
    np.testing.assert_allclose(home.info()[1, :10], away.info()[1, :10], rtol=1e-5, atol=1e-6)   # float atomics
    # the untouched neighbour slot of the importing population is still the placeholder it was
    fresh = Population(_cfg_dict(cfg), [1.0, 1.0], [0, 0])
    for k, v in _state(fresh, 0).items():
        np.testing.assert_array_equal(v, _state(away, 0)[k], err_msg=k)
    for pop in (home, away, fresh):
        pop.close()


def test_blob_of_another_network_shape_is_refused():
    from fqlpop.population import Population, lib
    cfg = orc.FQLConfig(ob_dim=28, action_dim=5, batch_size=64)
    a = Population(_cfg_dict(cfg), [3.0], [1])
    d = _cfg_dict(cfg)
    d["value_hidden_dims"] = (256,) * 4
    d["layer_norm"] = False
    b = Population(d, [3.0], [2])
    assert a.member_state_words() != b.member_state_words()
    blob = a.export_member(0)
    # same word count but a header that says otherwise: corrupt the layout check word
    bad = blob.clone()
    bad[7] = 0
    before = _state(a, 0)
    s = torch.cuda.current_stream().cuda_stream
    assert lib.fqlpop_member_import(a._h, 0, bad.data_ptr(), s) != 0
    assert "same network shapes" in lib.fqlpop_last_error().decode()
    assert lib.fqlpop_member_import(a._h, 5, blob.data_ptr(), s) != 0        # no such slot
    for k, v in _state(a, 0).items():
        np.testing.assert_array_equal(v, before[k], err_msg=k)
    a.close()
    b.close()
