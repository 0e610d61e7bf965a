"""Parity of world-model training (fqlpop_wm_* through the C-ABI) with the CPU oracle oracle/wm_oracle.py (GPU only).

The kernels compute in fp32 like the reference; the oracle runs in float64 with autograd. Tolerances:
  * logs (losses, metrics): rtol 2e-5 / atol 1e-6;
  * gradients of one step: max-abs error <= 2e-5 * max|g| per leaf (fp32 accumulation order);
  * parameters after a few Adam steps: atol 2e-4 at lr 1e-3 -- the first Adam steps move every weight by ~lr * sign(g),
    so a gradient whose sign is decided by rounding moves by up to lr in either direction; the FRACTION of such entries
    is bounded (< 0.1 %) and everything else must agree to 2e-5.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import fql_oracle as orc  # noqa: E402  (checker only)
from oracle import philox_oracle as phx  # noqa: E402
from oracle import wm_oracle as wmo  # noqa: E402

O, A, B = 28, 5, 256
HID = (128, 256, 128)
THID = (128, 256, 128)


def _t64(tree):
    return {k: (_t64(v) if isinstance(v, dict) else torch.as_tensor(np.asarray(v), dtype=torch.float64)) for k, v in tree.items()}


def _np(tree):
    return {k: (_np(v) if isinstance(v, dict) else v.detach().cpu().numpy().astype(np.float32)) for k, v in tree.items()}


def _batch(seed, rows=B, p_term=0.1):
    g = torch.Generator().manual_seed(seed)
    obs = torch.randn(rows, O, generator=g, dtype=torch.float64)
    act = torch.rand(rows, A, generator=g, dtype=torch.float64) * 2 - 1
    nxt = obs + 0.05 * torch.randn(rows, O, generator=g, dtype=torch.float64)
    rew = -(torch.rand(rows, generator=g) >= p_term).to(torch.float64)  # 0 with probability p_term, else -1
    return {"observations": obs, "actions": act, "next_observations": nxt, "rewards": rew}


def _f32(batch):
    return {k: v.to(torch.float32).numpy() for k, v in batch.items()}


def _envmodel(seed=5):
    em = orc.init_envmodel(seed, O, A, HID, dtype=torch.float64)
    g = torch.Generator().manual_seed(seed + 1)
    for net in ("state", "termination"):
        for name, leaf in orc.tree_leaves(em[net]):
            if name.endswith("bias") or name.endswith("scale"):
                leaf.add_(0.05 * torch.randn(leaf.shape, generator=g, dtype=torch.float64))
    return em


def _check_tree(got, want, rel, what):
    for (name, w) in wmo.tree_leaves(want):
        gv = got
        for k in name.split("/"):
            gv = gv[k]
        w = w.detach().numpy()
        err = np.abs(gv.astype(np.float64) - w).max()
        assert err <= rel * max(np.abs(w).max(), 1e-12), f"{what} {name}: max err {err:.3e} vs scale {np.abs(w).max():.3e}"


def _check_params(got, want, lr):
    tot = bad = 0
    for (name, w) in wmo.tree_leaves(want):
        gv = got
        for k in name.split("/"):
            gv = gv[k]
        d = np.abs(gv.astype(np.float64) - w.detach().numpy())
        assert d.max() <= 4 * lr, f"{name}: {d.max():.3e}"
        tot += d.size
        bad += int((d > 2e-5).sum())
    # Adam divides by sqrt(v): a parameter whose gradient sits at fp32 rounding level (|g| ~ 1e-10 against leaf scales of
    # 1e-3 in the smallest case, T=6 x 64 rows) moves by O(lr) in either direction, so a few per mille of the parameters
    # differ between ANY two fp32 summation orders (per-step launches vs scan kernels: 167 vs 73 of 73950 against the
    # float64 oracle, gradients equal to 1e-7 of their scale -- profiles/ubench/wm_scan_vs_steps.py); all stay within 4 lr
    assert bad <= 3e-3 * tot, f"{bad} of {tot} parameters differ by more than 2e-5"


def _logs_close(got, want, keys):
    for k in keys:
        assert got[k] == pytest.approx(want[k], rel=2e-5, abs=1e-6), (k, got[k], want[k])


@pytest.mark.parametrize("termination_weight", [1.0, 0.0], ids=["with_termination", "mse_only"])
def test_state_predictor_steps_match_oracle(termination_weight):
    from fqlpop.worldmodel import STATE, WorldModelTrainer
    em = _envmodel()
    params, term = em["state"], em["termination"]
    tr = WorldModelTrainer(STATE, O, A, HID, B, steps=100, init_learning_rate=1e-3, params=_np(params),
                           termination_params=_np(term) if termination_weight > 0 else None,
                           termination_hidden_dims=THID, true_termination_weight=30.0,
                           termination_weight=termination_weight)
    opt = wmo.new_opt_state(params)
    keys = ["loss", "next_observation_loss"] + (["termination_loss", "true_termination_loss", "false_termination_loss"]
                                                if termination_weight > 0 else [])
    for step in range(3):
        batch = _batch(10 + step)
        params, want, g = wmo.state_step(params, opt, batch, term, lr0=1e-3, decay_steps=100,
                                         true_termination_weight=30.0, termination_weight=termination_weight)
        got = tr.step(_f32(batch), train=True)
        _logs_close(got, want, keys)
        if step == 0:
            _check_tree(tr.grads(), g, 2e-5, "grad")
        assert tr.logs()[10] == pytest.approx(want["learning_rate"], rel=1e-5)
    _check_params(tr.params(), params, 1e-3)
    # eval_step: forward only, parameters untouched
    before = tr.export(0)
    batch = _batch(77)
    _, want, _ = wmo.state_step(_t64(tr.params()), None, batch, term, lr0=1e-3, decay_steps=100,
                                true_termination_weight=30.0, termination_weight=termination_weight, train=False)
    _logs_close(tr.step(_f32(batch), train=False), want, keys)
    assert np.array_equal(before, tr.export(0))
    assert tr.logs()[11] == 3.0


def test_termination_predictor_steps_match_oracle():
    from fqlpop.worldmodel import TERMINATION, WorldModelTrainer
    params = _envmodel(9)["termination"]
    tr = WorldModelTrainer(TERMINATION, O, A, THID, B, steps=50, init_learning_rate=1e-3, params=_np(params))
    opt = wmo.new_opt_state(params)
    g = torch.Generator().manual_seed(3)
    mask = (torch.rand(B, O, generator=g) >= 0.1).to(torch.float64) / 0.9  # nn.Dropout(0.1) multiplier, fixed
    for step in range(3):
        batch = _batch(20 + step)
        params, want, gr = wmo.termination_step(params, opt, batch, mask, lr0=1e-3, decay_steps=50)
        got = tr.step(_f32(batch), train=True, dropout_mask=mask.to(torch.float32).numpy())
        _logs_close(got, want, ["loss", "true_loss", "false_loss"])
        if step == 0:
            _check_tree(tr.grads(), gr, 2e-5, "grad")
    _check_params(tr.params(), params, 1e-3)
    batch = _batch(88, p_term=0.5)
    _, want, _ = wmo.termination_step(_t64(tr.params()), None, batch, lr0=1e-3, decay_steps=50, train=False)
    got = tr.step(_f32(batch), train=False)
    _logs_close(got, want, ["loss", "true_loss", "false_loss", "accuracy", "precision", "recall", "f1"])


def test_cosine_schedule_reaches_zero_and_stays():
    from fqlpop.worldmodel import TERMINATION, WorldModelTrainer
    tr = WorldModelTrainer(TERMINATION, O, A, (32,), B, steps=4, init_learning_rate=1e-2, seed=1)
    batch = _f32(_batch(1))
    lrs = []
    for _ in range(6):
        tr.step(batch, train=True)
        lrs.append(float(tr.logs()[10]))
    want = [wmo.cosine_lr(c, 1e-2, 4) for c in range(6)]
    assert lrs == pytest.approx(want, rel=1e-5, abs=1e-9)
    assert lrs[4] == pytest.approx(0.0, abs=1e-9) and lrs[5] == pytest.approx(0.0, abs=1e-9)


def test_resident_dataset_sampling_matches_explicit_batches():
    """fqlpop_wm_train draws row Philox(seed, step, b, 0) % N for batch row b of optimiser step `step`; feeding the same
    rows through fqlpop_wm_step must give bit-identical parameters."""
    from fqlpop.worldmodel import STATE, WorldModelTrainer
    em = _envmodel(11)
    data = _f32(_batch(5, rows=5000))
    kw = dict(steps=20, params=_np(em["state"]), termination_params=_np(em["termination"]), termination_hidden_dims=THID, seed=1234)
    a = WorldModelTrainer(STATE, O, A, HID, B, **kw)
    b = WorldModelTrainer(STATE, O, A, HID, B, **kw)
    a.set_dataset(data)
    a.train(4)
    for step in range(4):
        idx = np.array([phx.index(1234, step, r, 0, 5000) for r in range(B)])
        b.step({k: v[idx] for k, v in data.items()}, train=True)
    assert np.array_equal(a.export(0), b.export(0))
    assert a.logs()[11] == 4.0


def test_training_learns_a_synthetic_transition_function():
    """Sanity at scale: 300 device-sampled steps on next = obs + 0.1 * tanh(obs @ W) reduce the validation loss >5x."""
    from fqlpop.worldmodel import STATE, WorldModelTrainer
    g = torch.Generator().manual_seed(0)
    N = 20000
    obs = torch.randn(N, O, generator=g)
    act = torch.rand(N, A, generator=g) * 2 - 1
    W = torch.randn(O + A, O, generator=g) / 6
    nxt = obs + 0.1 * torch.tanh(torch.cat([obs, act], 1) @ W)
    data = {"observations": obs.numpy(), "actions": act.numpy(), "next_observations": nxt.numpy(),
            "rewards": -np.ones(N, np.float32)}
    tr = WorldModelTrainer(STATE, O, A, HID, B, steps=300, termination_weight=0.0, seed=3)
    val = {k: v[:B] for k, v in data.items()}
    first = tr.step(val, train=False)["next_observation_loss"]
    tr.set_dataset(data)
    tr.train(300)
    last = tr.step(val, train=False)["next_observation_loss"]
    assert np.isfinite(last) and last < first / 5, (first, last)


def test_trainers_train_save_and_reload(tmp_path):
    """The reference's two-stage recipe end to end on synthetic data: train the termination predictor, save it in the
    reference's file layout, train the state predictor against it (termination_weight > 0 loads it through load_model),
    save, and read both back with load_model for the simulated task."""
    from functools import partial
    from envmodel.baseline import BaselineStatePredictor, TerminationPredictor
    from envmodel.config import TrainerConfig
    from envmodel.loss import focal_loss, state_prediction_loss
    from envmodel.state_predictor_trainer import StatePredictorTrainer
    from envmodel.termination_predictor_trainer import TerminationPredictorTrainer
    from utils.data_loader import StepLoader
    from utils.envmodel import load_model, save_model

    g = torch.Generator().manual_seed(0)
    N = 8000
    obs = torch.randn(N, O, generator=g)
    act = torch.rand(N, A, generator=g) * 2 - 1
    nxt = obs + 0.1 * torch.tanh(torch.cat([obs, act], 1) @ (torch.randn(O + A, O, generator=g) / 6))
    rew = -(nxt[:, 0] < 1.0).float()  # "terminated" (reward 0) where the first coordinate is large: learnable, ~16 % positives
    data = {"observations": obs.numpy(), "actions": act.numpy(), "next_observations": nxt.numpy(), "rewards": rew.numpy(),
            "masks": (rew != 0).float().numpy(), "terminals": np.zeros(N, np.float32)}
    np.random.seed(0)
    logged = []

    class Log:
        def log(self, d, step):
            logged.append((step, d))

    cfg = TrainerConfig(steps=400, env_name="synthetic-env", model="termination_predictor", val_batches=2,
                        model_config={"hidden_dims": (64, 64)}, save_directory=tmp_path)
    tt = TerminationPredictorTrainer(TerminationPredictor(O, (64, 64)), StepLoader(data), StepLoader(data), focal_loss, cfg,
                                     logger=Log(), device_sampling=True)
    tt.train()
    first_val = [d["val/loss"] for st_, d in logged if st_ == 0 and "val/loss" in d][0]
    assert tt.last_validation["loss"] < first_val / 2 and tt.last_validation["accuracy"] > 0.84 and tt.state.step == 400
    assert any("val/f1" in d for _, d in logged) and any("train/learning_rate" in d for _, d in logged)
    save_model(tt.state.params, tmp_path, cfg.env_name, "termination_predictor", cfg.model_config)

    cfg2 = TrainerConfig(steps=150, env_name="synthetic-env", model="baseline", val_batches=2,
                         model_config={"hidden_dims": (64, 64)}, save_directory=tmp_path)
    loss = partial(state_prediction_loss, true_termination_weight=cfg2.true_termination_weight,
                   termination_weight=cfg2.termination_weight, reconstruction_weight=0.0)
    st = StatePredictorTrainer(BaselineStatePredictor(O, A, (64, 64)), StepLoader(data), StepLoader(data), loss, cfg2)
    first = st._validate(0)
    st.train()  # host-sampled batches: one np.random draw per step, as the reference
    assert st.last_validation["next_observation_loss"] < first["next_observation_loss"] / 2
    assert set(st.last_validation) == {"loss", "next_observation_loss", "termination_loss", "true_termination_loss",
                                      "false_termination_loss"}
    save_model(st.state.params, tmp_path, cfg2.env_name, "baseline", cfg2.model_config)

    sample = {k: v[:4] for k, v in data.items()}
    sp = load_model(tmp_path, "synthetic-env", "baseline", sample)
    tp = load_model(tmp_path, "synthetic-env", "termination_predictor", sample)
    pred, same = sp(sample["observations"], sample["actions"])
    assert pred.shape == (4, O) and np.array_equal(same, sample["observations"])
    want = orc.state_predictor_apply(_t64(st.state.params["params"]), torch.as_tensor(sample["observations"], dtype=torch.float64),
                                     torch.as_tensor(sample["actions"], dtype=torch.float64)).numpy()
    assert np.abs(pred - want).max() < 3e-2  # the rollout kernels multiply in bf16
    assert tp(sample["next_observations"]).shape == (4,)


def test_initial_observation_decoder_matches_oracle_and_loads_from_reference_files(tmp_path):
    """Decoder z -> [Dense, relu]* -> Dense(O) (envmodel/initial_observation.py:21-30): explicit noise against a float64
    restatement (fp32 kernels: rtol 1e-5), device Philox noise against oracle/philox_oracle.py, ragged sample counts, and
    the reference's file layout (initial_observation.pt holds encoder AND decoder; hidden_dims is stored encoder-order)."""
    import yaml
    from envmodel.initial_observation import load_initial_observation_decoder
    from utils.flax_msgpack import to_bytes
    latent, enc_hidden = 8, (128, 64)
    dims = [latent, *enc_hidden[::-1], O]
    g = torch.Generator().manual_seed(4)
    dec = {f"Dense_{i}": {"kernel": torch.randn(dims[i], dims[i + 1], generator=g, dtype=torch.float64) / np.sqrt(dims[i]),
                          "bias": 0.1 * torch.randn(dims[i + 1], generator=g, dtype=torch.float64)} for i in range(3)}
    root = tmp_path / "synthetic-env" / "env_models"
    root.mkdir(parents=True)
    with open(root / "initial_observation_config.yaml", "w") as f:
        yaml.dump({"hidden_dims": enc_hidden, "latent_dim": latent}, f)
    with pytest.raises(FileNotFoundError, match="Model file not found"):
        load_initial_observation_decoder(tmp_path, "synthetic-env", O)
    (root / "initial_observation.pt").write_bytes(to_bytes({"params": {"decoder": _np(dec), "encoder": {"Dense_0": {
        "kernel": np.zeros((O, 4), np.float32), "bias": np.zeros(4, np.float32)}}}}))
    d = load_initial_observation_decoder(tmp_path, "synthetic-env", O)
    assert d.hidden_dims == enc_hidden[::-1]

    def ref(z):
        x = z
        for i in range(3):
            x = x @ dec[f"Dense_{i}"]["kernel"] + dec[f"Dense_{i}"]["bias"]
            if i < 2:
                x = torch.relu(x)
        return x.numpy()

    for n in (1, 33, 1100):   # 1100 = the sample count of distribution_visualizer.py:97
        z = torch.randn(n, latent, generator=g, dtype=torch.float64)
        got = d.sample(n, noise=z.to(torch.float32)).cpu().numpy()
        np.testing.assert_allclose(got, ref(z), rtol=1e-5, atol=1e-5)
    n = 64
    z = torch.tensor([[phx.normal(77, 0, r, c) for c in range(latent)] for r in range(n)], dtype=torch.float64)
    np.testing.assert_allclose(d.sample(n, seed=77).cpu().numpy(), ref(z), rtol=1e-4, atol=1e-4)
    a, b = d.sample(n, seed=78), d.sample(n, seed=78)
    assert torch.equal(a, b) and not torch.equal(a, d.sample(n, seed=79))


@pytest.mark.parametrize("T,rows", [(6, 64), (40, 256), (5, 30), (37, 250)],
                         ids=["T6_B64", "T40_B256", "T5_B30_ragged_row_tile", "T37_B250_ragged_rows_in_the_64x64_gemm"])
def test_multistep_predictor_steps_match_oracle(T, rows):
    """Back-propagation through time of the scanned cell (envmodel/multistep.py) against torch autograd in float64:
    logs, every gradient leaf of the first step, parameters after three Adam steps, and a forward-only evaluation."""
    from fqlpop.worldmodel import MULTISTEP, WorldModelTrainer
    em = _envmodel(21)
    # a trained cell predicts small residuals; keep the synthetic one contractive so T steps stay O(1)
    em["state"][f"Dense_{len(HID)}"]["kernel"].mul_(0.1)
    params, term = em["state"], em["termination"]
    tr = WorldModelTrainer(MULTISTEP, O, A, HID, rows, steps=100, params=_np(params), termination_params=_np(term),
                           termination_hidden_dims=THID, sequence_length=T)
    opt = wmo.new_opt_state(params)
    keys = ["loss", "next_observation_loss", "termination_loss", "true_termination_loss", "false_termination_loss"]

    def window(seed):
        g = torch.Generator().manual_seed(seed)
        obs = torch.randn(rows, T, O, generator=g, dtype=torch.float64)
        return {"observations": obs, "actions": torch.rand(rows, T, A, generator=g, dtype=torch.float64) * 2 - 1,
                "next_observations": obs + 0.05 * torch.randn(rows, T, O, generator=g, dtype=torch.float64),
                "rewards": -(torch.rand(rows, T, generator=g) >= 0.1).to(torch.float64)}

    for step in range(3):
        batch = window(30 + step)
        params, want, g = wmo.multistep_step(params, opt, batch, term, lr0=1e-3, decay_steps=100)
        got = tr.step(_f32(batch), train=True)
        _logs_close(got, want, keys)
        if step == 0:  # fp32 sums over T*B rows through T chained steps: 5e-5 of the leaf's scale at T=6, 3e-4 at T=40
            _check_tree(tr.grads(), g, 5e-5 if T <= 8 else 3e-4, "grad")
    _check_params(tr.params(), params, 1e-3)
    batch = window(99)
    _, want, _ = wmo.multistep_step(_t64(tr.params()), None, batch, term, lr0=1e-3, decay_steps=100, train=False)
    _logs_close(tr.step(_f32(batch), train=False), want, keys)


def test_multistep_trainer_checkpoint_loads_as_rollout_model(tmp_path):
    """StatePredictorTrainer with the MultistepStatePredictor on MultistepLoader windows (1000-step episodes): the loss
    falls, and the checkpoint (params/ScanCell_0/cell, the reference's nesting) loads through load_model("multistep")."""
    from functools import partial
    from envmodel.config import TrainerConfig
    from envmodel.loss import state_prediction_loss
    from envmodel.multistep import MultistepStatePredictor
    from envmodel.state_predictor_trainer import StatePredictorTrainer
    from utils.data_loader import MultistepLoader
    from utils.envmodel import load_model, save_model
    g = torch.Generator().manual_seed(1)
    N = 4000  # four episodes
    act = torch.rand(N, A, generator=g) * 2 - 1
    W = torch.randn(O + A, O, generator=g) / 8
    obs = torch.zeros(N, O)
    nxt = torch.zeros(N, O)
    for ep in range(4):
        o = 0.5 * torch.randn(O, generator=g)
        for i in range(ep * 1000, (ep + 1) * 1000):
            obs[i] = o
            o = 0.9 * o + 0.1 * torch.tanh(torch.cat([o, act[i]]) @ W)
            nxt[i] = o
    data = {"observations": obs.numpy(), "actions": act.numpy(), "next_observations": nxt.numpy(),
            "rewards": -np.ones(N, np.float32)}
    np.random.seed(2)
    cfg = TrainerConfig(steps=120, env_name="synthetic-env", model="multistep", val_batches=1, batch_size=32,
                        sequence_length=8, model_config={"hidden_dims": (64, 64)}, save_directory=tmp_path)
    loss = partial(state_prediction_loss, true_termination_weight=30.0, termination_weight=0.0, reconstruction_weight=0.0)
    st = StatePredictorTrainer(MultistepStatePredictor(O, A, (64, 64)), MultistepLoader(data, sequence_length=8),
                               MultistepLoader(data, sequence_length=8), loss, cfg)
    first = st._validate(0)["next_observation_loss"]
    st.train()
    assert st.last_validation["next_observation_loss"] < first / 3
    tree = st.state.params
    assert set(tree["params"]) == {"ScanCell_0"} and "Dense_0" in tree["params"]["ScanCell_0"]["cell"]
    save_model(tree, tmp_path, "synthetic-env", "multistep", cfg.model_config)
    sp = load_model(tmp_path, "synthetic-env", "multistep", {k: v[:4] for k, v in data.items()})
    pred, _ = sp(data["observations"][:4], data["actions"][:4])
    want = orc.state_predictor_apply(_t64(tree["params"]["ScanCell_0"]["cell"]),
                                     torch.as_tensor(data["observations"][:4], dtype=torch.float64),
                                     torch.as_tensor(data["actions"][:4], dtype=torch.float64)).numpy()
    assert np.abs(pred - want).max() < 3e-2  # the rollout kernels multiply in bf16


def test_resident_window_sampling_matches_explicit_windows():
    """fqlpop_wm_train for the multistep predictor: batch row b of optimiser step s takes episode Philox(seed, s, b, 0) %
    n_episodes and window start Philox(seed, s, b, 1) % (1000 - T); the same windows fed explicitly give identical bits."""
    from fqlpop.worldmodel import MULTISTEP, WorldModelTrainer
    T, rows, n_ep = 12, 32, 3
    em = _envmodel(31)
    em["state"][f"Dense_{len(HID)}"]["kernel"].mul_(0.1)
    data = _f32(_batch(8, rows=n_ep * 1000))
    kw = dict(steps=20, params=_np(em["state"]), termination_params=_np(em["termination"]), termination_hidden_dims=THID,
              seed=77, sequence_length=T)
    a = WorldModelTrainer(MULTISTEP, O, A, HID, rows, **kw)
    b = WorldModelTrainer(MULTISTEP, O, A, HID, rows, **kw)
    a.set_dataset(data)
    a.train(3)
    for step in range(3):
        idx = np.stack([phx.index(77, step, r, 0, n_ep) * 1000 + phx.index(77, step, r, 1, 1000 - T) + np.arange(T)
                        for r in range(rows)])                       # [rows, T]
        b.step({k: v[idx] for k, v in data.items()}, train=True)
    assert np.array_equal(a.export(0), b.export(0))
    assert a.logs()[11] == 3.0


@pytest.mark.parametrize("latent,hidden,rw,rows", [(8, (128, 64), 1.0, 256), (4, (128,), 10.0, 64), (3, (33, 50, 20), 0.0, 37)],
                         ids=["two_hidden", "reference_default_one_hidden", "odd_widths_ragged_batch_kl_only"])
def test_vae_steps_match_oracle(latent, hidden, rw, rows):
    """InitialObservationEnvModel + vae_loss (envmodel/initial_observation.py:32-78) with the backward pass and Adam on
    the device, against the float64 torch restatement fed the same batches and the same reparameterisation noise:
    logs to 2e-5, gradients of the first step to 2e-5 of each leaf's scale, parameters after 3 steps as the other
    world-model trainers; then an eval step leaves the parameters untouched."""
    from fqlpop.worldmodel import VaeTrainer, init_vae_params
    p0 = init_vae_params(3, O, latent, hidden)
    g = torch.Generator().manual_seed(9)
    for part in p0.values():                      # non-zero biases so that their gradients matter
        for layer in part.values():
            layer["bias"] = (0.05 * torch.randn(layer["bias"].shape, generator=g)).numpy().astype(np.float32)
    tr = VaeTrainer(O, latent, hidden, rows, steps=100, init_learning_rate=1e-3, reconstruction_weight=rw, params=p0)
    params = _t64(p0)
    opt = wmo.new_opt_state(params)
    for step in range(3):
        obs = torch.randn(rows, O, generator=g, dtype=torch.float64)
        eps = torch.randn(rows, latent, generator=g, dtype=torch.float64)
        params, want, grads = wmo.vae_step(params, opt, obs, eps, reconstruction_weight=rw, lr0=1e-3, decay_steps=100)
        got = tr.step(obs.to(torch.float32), eps.to(torch.float32), train=True)
        _logs_close(got, want, ["loss", "reconstruction_loss", "kl"])
        if step == 0:
            _check_tree(tr.grads(), grads, 2e-5, "grad")
        assert tr.logs()[10] == pytest.approx(want["learning_rate"], rel=1e-5)
    _check_params(tr.params(), params, 1e-3)
    before = tr.export(0)
    obs = torch.randn(rows, O, generator=g, dtype=torch.float64)
    eps = torch.randn(rows, latent, generator=g, dtype=torch.float64)
    _, want, _ = wmo.vae_step(params, opt, obs, eps, reconstruction_weight=rw, lr0=1e-3, decay_steps=100, train=False)
    _logs_close(tr.step(obs.to(torch.float32), eps.to(torch.float32), train=False), want, ["loss", "reconstruction_loss", "kl"])
    assert np.array_equal(before, tr.export(0)) and tr.logs()[11] == 3
    tr.close()


def test_vae_trains_on_the_device_and_its_decoder_generates_initial_observations(tmp_path):
    """train_env_model's `initial_observation` path end to end on synthetic episode starts drawn from a 2-dimensional
    manifold: resident dataset, batches and latent noise sampled on the device, loss falling, the files written where the
    reference's analysis script reads them, and the saved decoder (loaded through load_initial_observation_decoder, i.e.
    the rollout's source of initial observations) reproducing what the trainer's own decoder parameters give."""
    from functools import partial
    from envmodel.config import TrainerConfig
    from envmodel.initial_observation import (InitialObservationEnvModel, InitialObservationTrainer,
                                              load_initial_observation_decoder, vae_loss)
    from utils.data_loader import InitialObservationLoader
    from utils.envmodel import save_model
    rng = np.random.default_rng(0)
    n_eps = 300
    basis = rng.standard_normal((2, O)).astype(np.float32)
    starts = (rng.standard_normal((n_eps, 2)).astype(np.float32) @ basis + 0.01 * rng.standard_normal((n_eps, O))).astype(np.float32)
    obs = np.repeat(starts, 1000, axis=0)          # InitialObservationLoader keeps every 1000th row
    obs[np.arange(len(obs)) % 1000 != 0] += 5.0    # (rows that are not episode starts must not be sampled)
    loader = InitialObservationLoader({"observations": obs})
    assert loader.dataset["observations"].shape == (n_eps, O)
    cfg = TrainerConfig(env_name="synthetic-env", model="initial_observation", model_config={"hidden_dims": (64, 32), "latent_dim": 4},
                        steps=600, batch_size=128, init_learning_rate=3e-3, reconstruction_weight=50.0, val_batches=2,
                        save_directory=tmp_path)
    model = InitialObservationEnvModel(O, 4, (64, 32))
    tr = InitialObservationTrainer(model, loader, loader, partial(vae_loss, reconstruction_weight=50.0), cfg,
                                   device_sampling=True)
    first = tr.eval_step(loader.sample(128))
    tr.train()
    last = tr.last_validation
    assert tr.state.step == 600
    assert last["reconstruction_loss"] < 0.2 * first["reconstruction_loss"], (first, last)
    assert np.isfinite(list(last.values())).all()
    save_model(tr.state.params, tmp_path, "synthetic-env", "initial_observation", cfg.model_config)
    dec = load_initial_observation_decoder(tmp_path, "synthetic-env", O)
    z = torch.randn(50, 4, generator=torch.Generator().manual_seed(1))
    want = wmo.vae_decoder_apply(_t64(tr.state.params["params"]["decoder"]), z.to(torch.float64)).numpy()
    np.testing.assert_allclose(dec.sample(50, noise=z).cpu().numpy(), want, rtol=1e-4, atol=1e-4)
    # samples land near the data's 2-dimensional subspace: the residual off the span of `basis` is small against the data scale
    gen = dec.sample(512, seed=3).cpu().numpy()
    q, _ = np.linalg.qr(basis.T)
    off = gen - (gen @ q) @ q.T
    assert np.sqrt((off ** 2).mean()) < 0.35 * np.sqrt((starts ** 2).mean())
