"""The reference-facing host interfaces running on the real device population (B200): simulated task + world-model
files in the reference's on-disk format, Trainer with successive halving, single-experiment mode, agent save / load."""
import os
import pickle

import numpy as np
import pytest
import torch
import yaml

from oracle import fql_oracle as orc  # checker only

pytestmark = pytest.mark.gpu
ENV = "cube-single-play-singletask-task2-v0"


@pytest.fixture()
def workdir(tmp_path, monkeypatch):
    """<save>/<env>/env_models/{baseline,termination_predictor}{_config.yaml,.pt} as train_env_model.py writes them."""
    import fql.envs.env_utils as eu
    from utils import flax_msgpack
    monkeypatch.setattr(eu, "SYNTHETIC_ROWS", 20000)
    env = orc.init_envmodel(11, 28, 5, hidden=(128, 256, 128))
    env["termination"]["Dense_3"]["bias"] -= 0.5
    root = tmp_path / ENV / "env_models"
    root.mkdir(parents=True)
    to_np = lambda t: orc.tree_map(lambda x: x.numpy(), t)
    for name, key in (("baseline", "state"), ("termination_predictor", "termination")):
        (root / f"{name}.pt").write_bytes(flax_msgpack.to_bytes({"params": to_np(env[key])}))
        with open(root / f"{name}_config.yaml", "w") as f:
            yaml.dump({"hidden_dims": (128, 256, 128)}, f)
    return tmp_path, env


def _task(save_dir, n_envs=64, max_steps=30):
    from task.offline_task_simulated import OfflineTaskWithSimulatedEvaluations
    return OfflineTaskWithSimulatedEvaluations(ENV, "baseline", data_directory=None, save_directory=save_dir,
                                               buffer_size=1000, num_evaluation_envs=n_envs, max_episode_steps=max_steps)


def test_task_step_matches_oracle_and_reference_bookkeeping(workdir):
    save_dir, env = workdir
    task = _task(save_dir, n_envs=40, max_steps=3)
    obs, infos = task.reset(seed=5)
    assert obs.shape == (40, 28) and len(infos) == 40
    cur = torch.as_tensor(obs)
    seen_done = np.zeros(40, bool)
    for t in range(4):
        act = np.random.default_rng(t).uniform(-1, 1, (40, 5)).astype(np.float32)
        nxt, rew, term, trunc, infos = task.step(act)
        want = orc.state_predictor_apply(env["state"], cur, torch.as_tensor(act), emulate_bf16=True)
        np.testing.assert_allclose(nxt, want.numpy(), atol=3e-2, rtol=3e-2)
        assert np.array_equal(rew, np.where(term, 0, -1)) and trunc.all() == (t + 1 >= 3)
        assert [i.get("invalid", False) for i in infos] == seen_done.tolist()   # flagged from the step AFTER the first done
        seen_done |= term | trunc
        cur = torch.as_tensor(nxt)
    # load_model's stand-alone callables (reference API): state predictor returns (next, obs); termination returns logits
    nxt2, same = task.state_predictor(obs, act)
    assert np.array_equal(same, obs) and nxt2.shape == obs.shape
    logits = task.termination_predictor(obs, train=False, rng=None)
    want = orc.termination_predictor_apply(env["termination"], torch.as_tensor(obs), emulate_bf16=True)
    np.testing.assert_allclose(logits, want.numpy(), atol=3e-2, rtol=3e-2)


def test_trainer_successive_halving_on_device(workdir):
    from hpo.successive_halving import SuccessiveHalving
    from trainer.config import AgentConfig, ExperimentConfig, TrainerConfig
    from trainer.trainer import Trainer
    save_dir, _ = workdir
    task = _task(save_dir)
    cfg = TrainerConfig(steps=12, eval_interval=4, log_interval=2, save_directory=save_dir / "runs", env_name=ENV,
                        agent=AgentConfig(batch_size=64, actor_hidden_dims=(128, 128), value_hidden_dims=(512, 512)))
    pop = [ExperimentConfig(seed=s, alpha=a) for s, a in enumerate(np.logspace(np.log10(3), 3, 4).tolist())]
    sh = SuccessiveHalving(set(pop), total_evaluations=3, fraction=0.5, history_length=1)
    tr = Trainer(task, sh, cfg)
    tr.train(max_evaluations=50)
    steps = sorted(e.current_step for e in tr.experiments.values())
    assert steps[-1] == 12 and steps[0] == 4 and len(tr.candidates) in (1, 2)
    adam_steps, _ = tr.population.get_counters()
    assert sorted(adam_steps.tolist()) == steps                     # pruned members really stopped stepping
    for c, e in tr.experiments.items():
        root = save_dir / "runs" / ENV / e.experiment_name
        rows = open(root / "train.csv").read().strip().splitlines()
        assert rows[0].startswith("critic/critic_loss,") and len(rows) == 1 + e.current_step // 2
        vals = np.array([[float(x) for x in r.split(",")] for r in rows[1:]])
        assert np.isfinite(vals).all()
        assert len(open(root / "eval.csv").read().strip().splitlines()) == 1 + e.current_step // 4
    state = pickle.loads(pickle.dumps({"trainer": tr.state_dict(), "strategy": sh.state_dict()}))
    assert set(state["trainer"]["experiments"]) == set(pop)
    tr.population.close()


def test_single_experiment_save_and_load_agent(workdir):
    from evaluator.evaluation import evaluate_agent
    from trainer.config import AgentConfig, ExperimentConfig, TrainerConfig
    from trainer.experiment import Experiment
    from utils.agent import load_agent
    save_dir, _ = workdir
    task = _task(save_dir, n_envs=32, max_steps=8)
    cfg = TrainerConfig(steps=6, eval_interval=3, log_interval=3, save_directory=save_dir / "single", env_name=ENV,
                        agent=AgentConfig(batch_size=64))
    exp = Experiment(task, cfg, ExperimentConfig(seed=7, alpha=30.0))
    assert exp.agent.config["alpha"] == 30.0 and exp.agent.config["seed"] == 7 and exp.agent.config["action_dim"] == 5
    done = False
    while not done:                                     # tune_alpha.py --single_experiment loop
        done = exp.train(cfg.eval_interval)
        exp.save_agent(checkpoint=True)
    score = exp.evaluate()
    assert 0.0 <= score <= 1.0 and exp.current_step == 6
    exp.save_agent()
    root = save_dir / "single" / ENV / exp.experiment_name
    assert (root / "checkpoint_3.pkl").exists() and (root / "checkpoint_6.pkl").exists()
    state = pickle.load(open(root / "params.pkl", "rb"))["agent"]
    assert state["network"]["opt_state"]["0"]["count"] == 6
    k = state["network"]["params"]["modules_critic"]["value_net"]["Dense_0"]["kernel"]
    assert k.shape == (2, 33, 512)
    # host step-by-step evaluation (the reference's loop) also runs against the device task
    stats, transitions = evaluate_agent.__globals__["evaluate_actor_fn"](
        lambda observations, temperature: exp.agent.sample_actions(observations, seed=3), task, seed=1)
    assert 0.0 <= stats["success"] <= 1.0 and 1 <= len(transitions) <= 8
    # load into a fresh population: identical policy output on fixed noise
    obs = task.sample("train", 16)
    agent2 = load_agent(root, obs)
    z = np.random.default_rng(0).standard_normal((16, 5)).astype(np.float32)
    a1 = exp.agent.population.sample_actions(0, obs["observations"], z).cpu().numpy()
    a2 = agent2.population.sample_actions(0, obs["observations"], z).cpu().numpy()
    np.testing.assert_array_equal(a1, a2)
    exp.stop()
    exp.agent.close()
    agent2.close()


def test_resumed_experiment_continues_its_batch_and_noise_stream(workdir):
    """A member restored from its state dict draws, at its next update, exactly the dataset rows (and noise) the
    uninterrupted run draws: the per-member Philox step counter travels with the checkpoint."""
    from trainer.config import AgentConfig, ExperimentConfig, TrainerConfig
    from trainer.experiment import Experiment
    save_dir, _ = workdir
    task = _task(save_dir, n_envs=8, max_steps=4)
    cfg = TrainerConfig(steps=20, eval_interval=5, log_interval=100, save_directory=save_dir / "resume", env_name=ENV,
                        agent=AgentConfig(batch_size=64))
    a = Experiment(task, cfg, ExperimentConfig(seed=11, alpha=30.0))
    a.train(3)
    state = pickle.loads(pickle.dumps(a.state_dict()))
    assert state["agent"]["fqlpop"]["philox_step"] == 3
    b = Experiment(task, cfg, ExperimentConfig(seed=11, alpha=30.0), state_dict=state)
    assert int(b.agent.population.get_counters()[1][0]) == 3
    a.train(1)
    rew_a, mask_a = a.agent.population.last_batch()
    b._bound = None
    b.train(1)
    rew_b, mask_b = b.agent.population.last_batch()
    np.testing.assert_array_equal(rew_a, rew_b)
    np.testing.assert_array_equal(mask_a, mask_b)
    ia, ib = a.agent.population.info(), b.agent.population.info()
    np.testing.assert_allclose(ia[0, :10], ib[0, :10], rtol=1e-4, atol=1e-5)   # same rows, same noise, same parameters
    for e in (a, b):
        e.stop()
        e.agent.close()


def test_scripts_train_env_models_then_tune_alpha(tmp_path, monkeypatch):
    """The reference's workflow with its own command lines: train_env_model.py (termination predictor, then the state
    predictor against it) leaves the env-model files, tune_alpha.py sweeps alpha with evaluations in that world model and
    leaves checkpoint.pkl + per-experiment logs."""
    import fql.envs.env_utils as eu
    monkeypatch.setattr(eu, "SYNTHETIC_ROWS", 8000)
    import train_env_model
    import tune_alpha
    common = ["--env_name", ENV, "--save_directory", str(tmp_path), "--data_directory", str(tmp_path / "nodata")]
    wm = ["--steps", "60", "--val_batches", "2", "--model.hidden_dims", "(64, 64)"] + common
    t = train_env_model.main(["--model", "termination_predictor", "--device_sampling"] + wm)
    assert t.state.step == 60
    s = train_env_model.main(["--model", "baseline"] + wm)
    assert set(s.last_validation) >= {"loss", "next_observation_loss", "termination_loss"}
    root = tmp_path / ENV / "env_models"
    assert sorted(p.name for p in root.iterdir()) == ["baseline.pt", "baseline_config.yaml", "termination_predictor.pt",
                                                      "termination_predictor_config.yaml"]
    # the initial-observation VAE on the episode starts; the sweep below starts its simulated episodes from its samples
    v = train_env_model.main(["--model", "initial_observation", "--device_sampling", "--model.latent_dim", "4",
                              "--reconstruction_weight", "10.0"] + wm)
    assert v.state.step == 60 and set(v.last_validation) == {"loss", "reconstruction_loss", "kl"}
    assert (root / "initial_observation.pt").exists() and (root / "initial_observation_config.yaml").exists()
    trainer = tune_alpha.main(["--number_of_alphas", "3", "--number_of_seeds", "1", "--max_evaluations", "6",
                               "--steps", "40", "--eval_interval", "20", "--log_interval", "20", "--eval_episodes", "32",
                               "--agent.layer_norm", "--simulated_evaluations", "--max_episode_steps", "20",
                               "--generated_initial_observations",
                               "--strategy", "successive_halving", "--total_evaluations", "6", "--history_length", "1"]
                              + common)
    assert trainer.task.initial_observation_decoder is not None
    with open(tmp_path / ENV / "checkpoint.pkl", "rb") as f:
        ck = pickle.load(f)
    assert set(ck) == {"trainer", "strategy"}
    assert trainer.strategy.performed_evaluations == 2  # steps 40 / eval_interval 20
    # reproduce.py: seed sweep at the default alpha, single-experiment mode (train -> evaluate -> save_agent)
    import reproduce
    exp = reproduce.main(["--number_of_seeds", "2", "--single_experiment", "--job_id", "1", "--steps", "20",
                          "--eval_interval", "10", "--log_interval", "10", "--eval_episodes", "16", "--agent.layer_norm",
                          "--simulated_evaluations", "--max_episode_steps", "10"] + common)
    assert exp.current_step == 20


def test_checkpoint_sweep_scores_many_agents_in_one_rollout(workdir):
    """evaluate_checkpoints (the measurement loop of evaluate_success_rate_correlation.py): three saved checkpoints of two
    experiments scored in one batched rollout equal the one-at-a-time evaluate_agent results with the same seed."""
    from evaluator.checkpoint_sweep import evaluate_checkpoints, find_experiment, success_rate_correlation
    from evaluator.evaluation import evaluate_agent
    from trainer.config import AgentConfig, ExperimentConfig, TrainerConfig
    from trainer.experiment import Experiment
    from utils.agent import load_agent
    save_dir, _ = workdir
    task = _task(save_dir, n_envs=64, max_steps=12)
    cfg = TrainerConfig(steps=6, eval_interval=3, log_interval=3, save_directory=save_dir / "ckpt", env_name=ENV,
                        agent=AgentConfig(batch_size=64))
    names = []
    for seed, alpha in ((1, 3.0), (2, 300.0)):
        exp = Experiment(task, cfg, ExperimentConfig(seed=seed, alpha=alpha))
        while not exp.train(cfg.eval_interval):
            exp.save_agent(checkpoint=True)
        exp.save_agent(checkpoint=True)
        names.append(exp.experiment_name)
        exp.stop()
    root = save_dir / "ckpt"
    found = find_experiment(root, ENV, 1, "3.0")
    assert found is not None and found.name == names[0]
    assert find_experiment(root, ENV, 9, "3.0") is None
    ckpts = [(root / ENV / names[0], "checkpoint_3"), (root / ENV / names[0], "checkpoint_6"),
             (root / ENV / names[1], "checkpoint_6")]
    batched = evaluate_checkpoints(task, ckpts, batch=2, seed=5)     # 2 slots -> two fused launches
    single = []
    for directory, filename in ckpts:
        agent = load_agent(directory, task.sample("train", 1), agent_filename=filename)
        single.append(evaluate_agent(agent, task, seed=5)[0]["success"])
    assert batched == pytest.approx(single, abs=1e-6)
    stats = success_rate_correlation([0.1, 0.4, 0.35, 0.8], [0.2, 0.3, 0.5, 0.9])
    assert stats["spearman_rho"] == pytest.approx(0.8) and stats["kendall_tau"] == pytest.approx(2 / 3)
    assert 0.9 < stats["pearson_r"] <= 1.0
