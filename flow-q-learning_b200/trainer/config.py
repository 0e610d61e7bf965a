"""Run configuration of the alpha sweep. Field names and defaults are the interface of the reference
(/root/reference/trainer/config.py:5-44): argparser fills `TrainerConfig` / `AgentConfig` by keyword, `Experiment` overrides
agent fields with the non-None fields of an `ExperimentConfig`, the logger dumps them to config.yaml, checkpoints pickle them.

Fix rather than behaviour change: the nested agent default is a `default_factory` -- the reference's
`agent: AgentConfig = AgentConfig()` raises on Python >= 3.11 (mutable default).
The two big classes are generated from field tables; construct them with keywords."""
from dataclasses import dataclass, field, make_dataclass
from pathlib import Path
from typing import Optional, Tuple

_MLP = (512, 512, 512, 512)

_AGENT_FIELDS = [
    # networks
    ("actor_hidden_dims", Tuple[int, ...], _MLP),
    ("value_hidden_dims", Tuple[int, ...], _MLP),
    ("layer_norm", bool, True),            # critic LayerNorm
    ("actor_layer_norm", bool, False),
    ("encoder", Optional[str], None),      # visual encoder name; unused for state observations
    ("ob_dims", Optional[int], None),      # filled by FQLAgent.create from the example batch
    ("action_dim", Optional[int], None),   # idem
    # optimisation
    ("lr", float, 3e-4),
    ("batch_size", int, 256),
    ("tau", float, 0.005),                 # Polyak rate of the target critic
    ("discount", float, 0.99),
    # FQL
    ("alpha", float, 10.0),                # BC coefficient -- the hyper-parameter the sweep tunes
    ("flow_steps", int, 10),
    ("q_agg", str, "mean"),                # mean | min over the critic ensemble in the TD target
    ("normalize_q_loss", bool, False),
    ("agent_name", str, "fql"),
    ("seed", int, 0),
]
AgentConfig = make_dataclass("AgentConfig", [(n, t, field(default=d)) for n, t, d in _AGENT_FIELDS], module=__name__)


@dataclass(frozen=True)
class ExperimentConfig:
    """Hashable key of one population member; non-None fields override AgentConfig by name."""
    seed: Optional[int] = None
    alpha: Optional[float] = None


_TRAINER_FIELDS = [
    ("env_name", str, field(default="cube-single-play-singletask-task2-v0")),
    ("agent", AgentConfig, field(default_factory=AgentConfig)),
    ("steps", int, field(default=1000000)),           # training steps per candidate
    ("eval_interval", int, field(default=100000)),    # steps between evaluations (= one strategy round)
    ("log_interval", int, field(default=5000)),
    ("eval_episodes", int, field(default=50)),
    ("buffer_size", int, field(default=2000000)),
    ("seed", int, field(default=0)),
    ("use_wandb", bool, field(default=False)),
    ("save_directory", Path, field(default=Path("exp/"))),
    ("data_directory", Path, field(default=Path("data/"))),
]
TrainerConfig = make_dataclass("TrainerConfig", _TRAINER_FIELDS, module=__name__)
