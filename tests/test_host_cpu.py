"""Host-side mirror of the reference interfaces, on the CPU: configs / argparser, data loaders, the flax-msgpack wire
format, CSV logging layout, and the Trainer's population loop (lockstep chunks, evaluation budget, pruning, resume)
driven through a fake device population."""
import pickle

import numpy as np
import pytest

from argparser import build_config_from_args, get_argparser
from fakes import FakePopulation, FakeTask
from hpo.identity import Identity
from hpo.successive_halving import SuccessiveHalving
from trainer.config import AgentConfig, ExperimentConfig, TrainerConfig
from trainer.trainer import Trainer
from utils import flax_msgpack
from utils.data_loader import InitialObservationLoader, MultistepLoader, StepLoader


def test_config_defaults_follow_the_reference():
    c = TrainerConfig()
    assert (c.steps, c.log_interval, c.eval_interval, c.eval_episodes, c.buffer_size) == (1000000, 5000, 100000, 50, 2000000)
    a = c.agent
    assert (a.lr, a.batch_size, a.discount, a.tau, a.q_agg, a.alpha, a.flow_steps) == (3e-4, 256, 0.99, 0.005, "mean", 10.0, 10)
    assert a.layer_norm and not a.actor_layer_norm and a.actor_hidden_dims == (512,) * 4
    assert TrainerConfig().agent is not c.agent                       # default_factory, not a shared instance
    assert hash(ExperimentConfig(1, 3.0)) == hash(ExperimentConfig(1, 3.0))


def test_argparser_namespacing_and_store_true_quirk():
    p = get_argparser()
    cfg = build_config_from_args(p.parse_args([]))
    assert cfg.agent.layer_norm is False            # store_true: False from the CLI although the dataclass says True
    cfg = build_config_from_args(p.parse_args(["--agent.layer_norm", "--agent.alpha", "300", "--eval_interval", "20000",
                                               "--agent.actor_hidden_dims", "(64, 64)", "--env_name", "antsoccer-x"]))
    assert cfg.agent.layer_norm and cfg.agent.alpha == 300.0 and cfg.eval_interval == 20000
    assert cfg.agent.actor_hidden_dims == (64, 64) and cfg.agent.value_hidden_dims == (512,) * 4
    assert cfg.env_name == "antsoccer-x" and cfg.save_directory.name == "exp"


def test_data_loaders_sampling_semantics():
    n = 3000
    ds = {"observations": np.arange(n * 2, dtype=np.float32).reshape(n, 2), "terminals": np.zeros(n, np.float32)}
    np.random.seed(0)
    b = StepLoader(dict(ds)).sample(16)
    np.random.seed(0)
    idx = np.random.randint(n, size=16)                                 # uniform with replacement, global numpy RNG
    np.testing.assert_array_equal(b["observations"], ds["observations"][idx])
    init = InitialObservationLoader(dict(ds)).sample(64)
    assert set((init["observations"][:, 0] / 2).astype(int)) <= {0, 1000, 2000}
    np.random.seed(1)
    w = MultistepLoader(dict(ds), sequence_length=8).sample(5)
    assert w["observations"].shape == (5, 8, 2) and w["terminals"].shape == (5, 8)
    rows = (w["observations"][:, :, 0] / 2).astype(int)
    assert np.all(np.diff(rows, axis=1) == 1) and np.all(rows[:, 0] // 1000 == rows[:, -1] // 1000)


def test_flax_msgpack_round_trip_and_layout():
    tree = {"params": {"Dense_0": {"kernel": np.random.randn(3, 4).astype(np.float32), "bias": np.zeros(4, np.float32)},
                       "LayerNorm_0": {"scale": np.ones(3, np.float32)}}, "step": np.int32(7)}
    blob = flax_msgpack.to_bytes(tree)
    back = flax_msgpack.from_bytes(blob)
    np.testing.assert_array_equal(back["params"]["Dense_0"]["kernel"], tree["params"]["Dense_0"]["kernel"])
    assert back["params"]["Dense_0"]["kernel"].dtype == np.float32 and int(back["step"]) == 7
    # the published flax layout: ExtType(1, packb((shape, dtype_name, bytes)))
    import msgpack
    raw = msgpack.unpackb(blob, raw=False)
    ext = raw["params"]["Dense_0"]["bias"]
    assert isinstance(ext, msgpack.ExtType) and ext.code == 1
    shape, dtype, buf = msgpack.unpackb(ext.data, raw=False)
    assert list(shape) == [4] and dtype == "float32" and len(buf) == 16


def _trainer(tmp_path, n=8, strategy="sh", steps=40, eval_interval=10, log_interval=5, total=4, state=None, task=None):
    cfg = TrainerConfig(steps=steps, eval_interval=eval_interval, log_interval=log_interval, save_directory=tmp_path,
                        env_name="cube-test", agent=AgentConfig(batch_size=8))
    pop = [ExperimentConfig(seed=i, alpha=a) for i, a in enumerate(np.logspace(np.log10(3), 3, n).tolist())]
    st = state or {}
    if strategy == "sh":
        s = SuccessiveHalving(set(pop), total_evaluations=total, fraction=0.5, history_length=1, state_dict=st.get("strategy"))
    else:
        s = Identity(pop, total_evaluations=0, state_dict=st.get("strategy"))
    task = task or FakeTask()
    return Trainer(task, s, cfg, state_dict=st.get("trainer"), population_factory=FakePopulation), pop, task


def test_trainer_lockstep_successive_halving(tmp_path):
    tr, pop, task = _trainer(tmp_path)
    tr.train(max_evaluations=100)
    fp = tr.population
    # milestones for N=8, total=4, f=.5, h=1 -> [2, 1, 1]: 8 -> 4 after eval 1, -> 2 after eval 2, then no more pruning
    assert sorted(fp.steps_done.tolist(), reverse=True)[:2] == [40, 40]
    assert sorted(fp.steps_done.tolist()) == [10, 10, 10, 10, 20, 20, 40, 40]
    best = sorted(pop, key=lambda c: abs(np.log10(c.alpha) - 2))[:2]
    assert set(tr.candidates) == set(best) and set(tr.finished_candidates) == set(best)
    # one fused evaluation per sweep, training issued for all active members at once, in chunks ending at log points
    assert task.rollouts == 4
    steps = [c for c in fp.calls if c[0] == "step"]
    assert all(c[1] == 5 for c in steps) and len(steps[0][2]) == 8
    # logs: eval.csv rows = evaluations done by that member; train.csv rows at every log_interval
    e = tr.experiments[best[0]]
    root = tmp_path / "cube-test" / e.experiment_name
    assert len(open(root / "eval.csv").read().strip().splitlines()) == 1 + 4
    assert len(open(root / "train.csv").read().strip().splitlines()) == 1 + 8
    assert open(root / "eval.csv").readline().strip() == "success,step"
    assert (root / "params.pkl").exists() and (root / "config.yaml").exists()


def test_trainer_budget_and_resume(tmp_path):
    tr, pop, _ = _trainer(tmp_path, n=4, strategy="identity")
    tr.train(max_evaluations=6)          # one full sweep (4) + 2 of the second
    assert sorted(tr.population.steps_done.tolist()) == [10, 10, 20, 20]
    assert len(tr.untrained_candidates) == 2
    state = pickle.loads(pickle.dumps({"trainer": tr.state_dict(), "strategy": tr.strategy.state_dict()}))
    tr2, _, _ = _trainer(tmp_path, n=4, strategy="identity", state=state)
    assert [tr2.experiments[c].current_step for c in tr2._ordered(tr2.experiments)] == [20, 20, 10, 10]
    tr2.train(max_evaluations=2)         # finishes the interrupted sweep first
    assert tr2.population.steps_done.tolist() == [20, 20, 20, 20] and tr2.untrained_candidates == []


def test_strategy_is_updated_by_config_not_by_experiment(tmp_path):
    tr, pop, _ = _trainer(tmp_path, n=4)
    tr.train(max_evaluations=4)
    assert set(tr.strategy.candidate_scores) == set(pop)


def test_missing_dataset_raises_unless_synthetic_is_opted_in(tmp_path, monkeypatch):
    """A wrong --data_directory must not silently train on noise (only FQLPOP_SYNTHETIC=1 generates data); the compact
    OGBench layout with a trailing non-terminal row loads without indexing past the end."""
    import numpy as np
    import pytest
    from fql.envs import env_utils
    monkeypatch.setenv("FQLPOP_SYNTHETIC", "0")
    with pytest.raises(FileNotFoundError):
        env_utils.make_env_and_datasets("cube-single-play-singletask-task2-v0", tmp_path)
    with pytest.raises(FileNotFoundError):
        env_utils.make_env_and_datasets("cube-single-play-singletask-task2-v0", None)
    rng = np.random.default_rng(0)
    n = 50
    term = np.zeros(n, np.float32); term[24] = 1.0                      # last row: terminals == 0
    np.savez(tmp_path / "cube-single-x.npz", observations=rng.standard_normal((n, 28)).astype(np.float32),
             actions=rng.uniform(-1, 1, (n, 5)).astype(np.float32), terminals=term)
    with pytest.raises(FileNotFoundError):                              # validation file missing, no opt-in
        env_utils.make_env_and_datasets("cube-single-x", tmp_path)
    monkeypatch.setenv("FQLPOP_SYNTHETIC", "1")
    monkeypatch.setattr(env_utils, "SYNTHETIC_ROWS", 2000)
    _, _, train, val = env_utils.make_env_and_datasets("cube-single-x", tmp_path)
    assert len(train["observations"]) == n - 2 and train["next_observations"].shape == train["observations"].shape
    _, _, train, _ = env_utils.make_env_and_datasets("cube-single-y", None)
    assert len(train["observations"]) == 2000


def _planner(world, owner_of, free):
    """A Trainer shell holding only what `_plan_moves` reads (the plan is a pure function of replicated state)."""
    t = object.__new__(Trainer)
    t.world = world
    cands = list(owner_of)
    t._order = {c: i for i, c in enumerate(cands)}
    t._owner = dict(owner_of)
    t._free_count = list(free)
    return t


def test_survivor_migration_plan():
    """Trainer._plan_moves: evens out the live candidates per rank with as few moves as possible, never into a rank
    without a free slot, never when ranks differ by at most one, and deterministically (every rank computes the plan)."""
    held = lambda t, moves, cands: [sum(1 for c in cands if dict((m[0], m[2]) for m in moves).get(c, t._owner[c]) == r)
                                    for r in range(t.world)]
    # 8 survivors on 4 of 8 GPUs (the 8-member rung of config 4): four moves, one member per GPU afterwards
    owner = {f"c{i}": i % 4 for i in range(8)}
    t = _planner(8, owner, [4] * 8)
    moves = t._plan_moves(list(owner))
    assert len(moves) == 4 and held(t, moves, list(owner)) == [1] * 8
    assert all(src < 4 <= dst for _, src, dst in moves)
    assert moves == _planner(8, owner, [4] * 8)._plan_moves(list(reversed(list(owner))))     # order of the set is irrelevant
    # only the candidates passed in count: pruned ones stay where they are and do not weigh
    assert t._plan_moves(["c0", "c5"]) == []                                    # ranks 0 and 1: already one each
    assert t._plan_moves(["c0", "c4"]) == [("c4", 0, 1)]                        # both on rank 0
    # a difference of one is left alone (5 members on 2 ranks as 3 + 2)
    t = _planner(2, {"a": 0, "b": 0, "c": 0, "d": 1, "e": 1}, [2, 2])
    assert t._plan_moves(list("abcde")) == []
    # no free slot anywhere else: nothing moves; one free slot: exactly one move, into it
    t = _planner(2, {"a": 0, "b": 0, "c": 0, "d": 0}, [0, 0])
    assert t._plan_moves(list("abcd")) == []
    t = _planner(3, {"a": 0, "b": 0, "c": 0, "d": 0}, [0, 0, 1])
    assert t._plan_moves(list("abcd")) == [("d", 0, 2)]
    # a slot freed by a departure can be refilled later in the same plan, and the free counts are not mutated
    t = _planner(2, {"a": 0, "b": 0, "c": 0, "d": 0}, [0, 4])
    moves = t._plan_moves(list("abcd"))
    assert held(t, moves, list("abcd")) == [2, 2] and t._free_count == [0, 4]
    # the last candidates in slot order are the ones that leave
    assert [m[0] for m in moves] == ["d", "c"]
