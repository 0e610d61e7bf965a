"""Restore one trained agent from its experiment directory: `config.yaml` (the agent's hyper-parameters) decides the
network shapes, `<agent_filename><agent_extension>` is a pickle {"agent": state_dict} written by Experiment.save_agent.
Signature, file names and error messages follow the reference's load_agent (/root/reference/utils/agent.py:9-41).
`population` / `slot` (optional, this build) put the agent into a slot of an existing device population."""
import pickle
from pathlib import Path

import yaml

from fql.agents.fql import FQLAgent, from_state_dict


def _existing(path: Path, what: str, hint: str) -> Path:
    if path.exists():
        return path
    raise FileNotFoundError(f"{what} not found at {path}. Please ensure the {hint}.")


def load_agent(agent_directory, sample_batch, agent_filename: str = "params", agent_extension: str = ".pkl",
               population=None, slot: int = 0):
    root = Path(agent_directory)
    hyper = yaml.unsafe_load(_existing(root / "config.yaml", "Configuration file", "configuration has been set").read_text())
    blob = _existing(root / f"{agent_filename}{agent_extension}", "Checkpoint", "agent has been trained").read_bytes()
    fresh = FQLAgent.create(hyper["seed"], sample_batch["observations"], sample_batch["actions"], hyper,
                            population=population, slot=slot)
    return from_state_dict(fresh, pickle.loads(blob)["agent"])
