"""Run configuration dataclasses -- same field names, defaults and hashing behaviour as the reference
(/root/reference/trainer/config.py:5-44), so argparser / Trainer / Experiment / checkpoints line up.

Difference that is a fix, not a behaviour change: the nested AgentConfig default is built with
`default_factory`, because the reference's `agent: AgentConfig = AgentConfig()` raises on Python >= 3.11.
"""
from dataclasses import dataclass, field
from pathlib import Path
from typing import Optional, Tuple


@dataclass
class AgentConfig:
    seed: int = 0
    agent_name: str = "fql"
    ob_dims: Optional[int] = None      # filled by FQLAgent.create from the example batch
    action_dim: Optional[int] = None   # idem
    lr: float = 3e-4
    batch_size: int = 256
    actor_hidden_dims: Tuple[int, ...] = (512, 512, 512, 512)
    value_hidden_dims: Tuple[int, ...] = (512, 512, 512, 512)
    layer_norm: bool = True
    actor_layer_norm: bool = False
    discount: float = 0.99
    tau: float = 0.005
    q_agg: str = "mean"
    alpha: float = 10.0
    flow_steps: int = 10
    normalize_q_loss: bool = False
    encoder: Optional[str] = None


@dataclass(frozen=True)
class ExperimentConfig:
    """Hashable key of one population member; non-None fields override AgentConfig by name."""
    seed: Optional[int] = None
    alpha: Optional[float] = None


@dataclass
class TrainerConfig:
    seed: int = 0
    steps: int = 1000000
    log_interval: int = 5000
    eval_interval: int = 100000
    save_directory: Path = Path("exp/")
    data_directory: Path = Path("data/")
    use_wandb: bool = False
    env_name: str = "cube-single-play-singletask-task2-v0"
    agent: AgentConfig = field(default_factory=AgentConfig)
    eval_episodes: int = 50
    buffer_size: int = 2000000
