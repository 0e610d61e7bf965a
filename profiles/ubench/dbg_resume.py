import os, sys, pickle
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "flow-q-learning_b200")); sys.path.insert(0, ROOT)
os.environ["FQLPOP_SYNTHETIC"] = "1"; os.environ["FQLPOP_SYNTHETIC_ROWS"] = "20000"
import numpy as np, torch, yaml, tempfile
from pathlib import Path
from oracle import fql_oracle as orc
from utils import flax_msgpack
ENV = "cube-single-play-singletask-task2-v0"
tmp = Path(tempfile.mkdtemp())
env = orc.init_envmodel(11, 28, 5, hidden=(128, 256, 128))
root = tmp / ENV / "env_models"; root.mkdir(parents=True)
to_np = lambda t: orc.tree_map(lambda x: x.numpy(), t)
for name, key in (("baseline", "state"), ("termination_predictor", "termination")):
    (root / f"{name}.pt").write_bytes(flax_msgpack.to_bytes({"params": to_np(env[key])}))
    yaml.dump({"hidden_dims": (128, 256, 128)}, open(root / f"{name}_config.yaml", "w"))
from task.offline_task_simulated import OfflineTaskWithSimulatedEvaluations
from trainer.config import AgentConfig, ExperimentConfig, TrainerConfig
from trainer.experiment import Experiment
task = OfflineTaskWithSimulatedEvaluations(ENV, "baseline", data_directory=None, save_directory=tmp, buffer_size=1000, num_evaluation_envs=8, max_episode_steps=4)
cfg = TrainerConfig(steps=20, eval_interval=5, log_interval=100, save_directory=tmp / "resume", env_name=ENV, agent=AgentConfig(batch_size=64))
a = Experiment(task, cfg, ExperimentConfig(seed=11, alpha=30.0))
a.train(3)
print("a counters", a.agent.population.get_counters(), "rew sum", a.agent.population.last_batch()[0].sum())
state = pickle.loads(pickle.dumps(a.state_dict()))
b = Experiment(task, cfg, ExperimentConfig(seed=11, alpha=30.0), state_dict=state)
print("b counters", b.agent.population.get_counters(), "current_step", b.current_step, "active", b.agent.population.active)
a.train(1)
print("a after", a.agent.population.get_counters(), a.agent.population.last_batch()[0].sum(), a.agent.population.info()[0, :4])
r = b.train(1)
print("b train returned", r, "step", b.current_step)
print("b after", b.agent.population.get_counters(), b.agent.population.last_batch()[0].sum(), b.agent.population.info()[0, :4])
b.train(1)
print("b after2", b.agent.population.get_counters(), b.agent.population.last_batch()[0].sum(), b.agent.population.info()[0, :4])
