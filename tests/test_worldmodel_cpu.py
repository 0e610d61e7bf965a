"""CPU checks of the world-model training oracle and host helpers (no GPU compute)."""
import ctypes
import math

import numpy as np
import pytest
import torch

from oracle import wm_oracle as wmo


def test_losses_known_answers():
    """Hand-computed values of envmodel/loss.py:7-62 on a 4-element batch."""
    x = torch.tensor([0.0, 2.0, -1.0, 3.0], dtype=torch.float64)
    t = torch.tensor([True, False, False, True])
    ce = [math.log(2), 2 + math.log1p(math.exp(-2)), math.log1p(math.exp(-1)), math.log1p(math.exp(-3))]
    loss, logs = wmo.weighted_binary_cross_entropy(x, t, true_weight=30.0)
    assert float(loss) == pytest.approx((30 * ce[0] + ce[1] + ce[2] + 30 * ce[3]) / 31 / 4, rel=1e-12)
    assert float(logs["true_loss"]) == pytest.approx((ce[0] + ce[3]) / 2, rel=1e-12)
    assert float(logs["false_loss"]) == pytest.approx((ce[1] + ce[2]) / 2, rel=1e-12)
    p = [1 / (1 + math.exp(-v)) for v in x.tolist()]
    fl = (0.25 * (1 - p[0]) ** 2 * ce[0] + 0.75 * p[1] ** 2 * ce[1] + 0.75 * p[2] ** 2 * ce[2] + 0.25 * (1 - p[3]) ** 2 * ce[3]) / 4
    loss, logs = wmo.focal_loss(x, t)
    assert float(loss) == pytest.approx(fl, rel=1e-12)
    # the MEAN loss scaled by count / (count + 1e-8); in fp32 (as in jax) the 1e-8 vanishes next to a count >= 1
    assert float(logs["true_loss"]) == pytest.approx(fl, rel=1e-7)
    pred, nxt = torch.tensor([[1.0, 2.0]], dtype=torch.float64), torch.tensor([[0.0, 4.0]], dtype=torch.float64)
    total, logs = wmo.state_prediction_loss(pred, nxt, x[:1], t[:1], pred, pred, 30.0, 1.0, 0.0)
    assert float(logs["next_observation_loss"]) == pytest.approx(2.5)
    assert float(total) == pytest.approx((2.5 + 30 * ce[0] / 31) / 2, rel=1e-12)
    total, logs = wmo.state_prediction_loss(pred, nxt, None, None, pred, pred, 30.0, 0.0, 0.0)
    assert float(total) == pytest.approx(2.5) and "termination_loss" not in logs


def test_oracle_adam_cosine_matches_torch_optim():
    """The restated optax.adam(cosine_decay_schedule) equals torch.optim.Adam under a cosine LambdaLR (same published
    algorithm, independent implementation)."""
    g = torch.Generator().manual_seed(0)
    p0 = torch.randn(64, generator=g, dtype=torch.float64)
    T, lr0 = 7, 1e-2
    params, opt = {"w": p0.clone()}, wmo.new_opt_state({"w": p0})
    ref = p0.clone().requires_grad_(True)
    o = torch.optim.Adam([ref], lr=lr0, betas=(0.9, 0.999), eps=1e-8)
    sched = torch.optim.lr_scheduler.LambdaLR(o, lambda c: 0.5 * (1 + math.cos(math.pi * min(c, T) / T)))
    for step in range(10):
        grad = torch.randn(64, generator=g, dtype=torch.float64)
        lr = wmo._adam(params, {"w": grad}, opt, lr0, T)
        assert lr == pytest.approx(wmo.cosine_lr(step, lr0, T))
        ref.grad = grad.clone()
        o.step()
        sched.step()
        assert torch.allclose(params["w"], ref.detach(), rtol=1e-12, atol=1e-15)
    assert wmo.cosine_lr(T, lr0, T) == pytest.approx(0.0, abs=1e-18) and wmo.cosine_lr(T + 5, lr0, T) == pytest.approx(0.0, abs=1e-18)


def test_flat_layout_round_trip_and_count():
    from fqlpop.worldmodel import (STATE, TERMINATION, WmConfig, flatten_params, init_params, layer_dims,
                                   unflatten_params)
    from fqlpop._lib import lib
    for kind, hidden in ((STATE, (128, 256, 128)), (TERMINATION, (128, 128))):
        tree = init_params(0, kind, 42, 8, hidden)
        flat = flatten_params(tree, kind)
        c = WmConfig()
        c.kind, c.ob_dim, c.action_dim, c.batch, c.n_hidden = kind, 42, 8, 256, len(hidden)
        for i, h in enumerate(hidden):
            c.hidden[i] = h
        assert lib.fqlpop_wm_num_params(ctypes.byref(c)) == flat.size
        back = unflatten_params(flat, kind, layer_dims(kind, 42, 8, hidden))
        assert sorted(back) == sorted(tree)
        for k in tree:
            for leaf in tree[k]:
                assert np.array_equal(back[k][leaf], tree[k][leaf])
    # flax defaults: LeCun-normal kernels truncated at 2 sigma, zero biases, LayerNorm scale 1
    tree = init_params(1, STATE, 42, 8, (128, 256, 128))
    k = tree["Dense_1"]["kernel"]
    assert k.shape == (128, 256) and abs(float(k.std()) - math.sqrt(1 / 128)) < 0.05 * math.sqrt(1 / 128)
    assert float(np.abs(k).max()) <= 2 * math.sqrt(1 / 128) / 0.87962566103423978 + 1e-6
    assert not tree["Dense_0"]["bias"].any() and (tree["LayerNorm_0"]["scale"] == 1).all()


def test_vae_parameter_layout_and_oracle_loss():
    """Flat VAE layout <-> flax tree (the mean and logvar heads share one device layer), the parameter count the library
    reports, and the oracle's vae_loss against the closed forms of envmodel/initial_observation.py:66-71 on a case whose
    answer is known: with every parameter zero, mu = logvar = 0, the reconstruction is 0, so kl = 0 and the
    reconstruction loss is mean(obs^2)."""
    import torch
    from oracle import wm_oracle as wmo
    from envmodel.initial_observation import InitialObservationEnvModel, vae_loss
    from fqlpop._lib import lib
    from fqlpop.worldmodel import VaeConfig, flatten_vae_params, init_vae_params, unflatten_vae_params, vae_dims
    O, L, hidden = 28, 4, (128, 64)
    tree = init_vae_params(0, O, L, hidden)
    assert sorted(tree["encoder"]) == ["Dense_0", "Dense_1", "Dense_2", "Dense_3"]      # two hidden, mean, logvar
    assert tree["encoder"]["Dense_2"]["kernel"].shape == (64, L) == tree["encoder"]["Dense_3"]["kernel"].shape
    assert [tree["decoder"][f"Dense_{i}"]["kernel"].shape for i in range(3)] == [(L, 64), (64, 128), (128, O)]
    assert vae_dims(O, L, hidden) == ([O, 128, 64, 2 * L], [L, 64, 128, O])
    m = InitialObservationEnvModel(O, L, hidden)
    assert m.encoder.hidden_dims == hidden and m.decoder.hidden_dims == hidden[::-1] and m.decoder.output_dimension == O
    flat = flatten_vae_params(tree, len(hidden))
    c = VaeConfig()
    c.ob_dim, c.latent_dim, c.batch, c.n_hidden = O, L, 64, len(hidden)
    for i, h in enumerate(hidden):
        c.hidden[i] = h
    lib.fqlpop_vae_num_params.argtypes = [ctypes.POINTER(VaeConfig)]
    lib.fqlpop_vae_num_params.restype = ctypes.c_int64
    assert lib.fqlpop_vae_num_params(ctypes.byref(c)) == flat.size
    back = unflatten_vae_params(flat, O, L, hidden)
    for part in tree:
        for layer in tree[part]:
            for leaf in tree[part][layer]:
                assert np.array_equal(back[part][layer][leaf], tree[part][layer][leaf]), (part, layer, leaf)
    t64 = lambda t: {k: (t64(v) if isinstance(v, dict) else torch.as_tensor(v, dtype=torch.float64)) for k, v in t.items()}
    g = torch.Generator().manual_seed(0)
    obs, eps = torch.randn(16, O, generator=g, dtype=torch.float64), torch.randn(16, L, generator=g, dtype=torch.float64)
    zero = t64({p: {l: {k: np.zeros_like(v) for k, v in leaves.items()} for l, leaves in tree[p].items()} for p in tree})
    _, logs = wmo.vae_loss(zero, obs, eps, 3.0)
    assert float(logs["kl"]) == 0.0
    assert float(logs["reconstruction_loss"]) == pytest.approx(float((obs ** 2).mean()))
    assert float(logs["loss"]) == pytest.approx(3.0 * float((obs ** 2).mean()) / 4.0)
    # the host-side numpy vae_loss agrees with the oracle on a random model
    p = t64(tree)
    mu, logvar = wmo.vae_encoder_apply(p["encoder"], obs)
    recon = wmo.vae_decoder_apply(p["decoder"], mu + eps * torch.exp(0.5 * logvar))
    loss, logs = wmo.vae_loss(p, obs, eps, 2.0)
    hl, hlogs = vae_loss(obs.numpy(), recon.numpy(), mu.numpy(), logvar.numpy(), 2.0)
    assert hl == pytest.approx(float(loss), rel=1e-5) and hlogs["kl"] == pytest.approx(float(logs["kl"]), rel=1e-5)
    # one oracle step moves the parameters and reports the cosine schedule's first learning rate
    new, slog, grads = wmo.vae_step(p, wmo.new_opt_state(p), obs, eps, reconstruction_weight=2.0, lr0=1e-3, decay_steps=10)
    assert slog["learning_rate"] == pytest.approx(1e-3)
    assert float((new["decoder"]["Dense_2"]["bias"] - p["decoder"]["Dense_2"]["bias"]).abs().max()) == pytest.approx(1e-3, rel=1e-3)


def test_initial_observation_loader_errors_like_the_reference(tmp_path):
    """Missing files raise FileNotFoundError with the reference's messages (distribution_visualizer.py:68-75,84-91)."""
    from envmodel.initial_observation import load_initial_observation_decoder
    with pytest.raises(FileNotFoundError, match="Configuration file not found"):
        load_initial_observation_decoder(tmp_path, "some-env", 28)
