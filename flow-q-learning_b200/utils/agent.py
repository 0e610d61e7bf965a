"""load_agent(agent_directory, sample_batch, agent_filename="params", agent_extension=".pkl") as the reference
(/root/reference/utils/agent.py:9-41): `config.yaml` -> FQLAgent.create -> pickle {"agent": state_dict} -> restore.
`population` / `slot` (extra, optional) load the agent into a slot of an existing device population."""
import pickle
from pathlib import Path

import yaml

from fql.agents.fql import FQLAgent, from_state_dict


def load_agent(agent_directory, sample_batch, agent_filename: str = "params", agent_extension: str = ".pkl",
               population=None, slot: int = 0):
    agent_directory = Path(agent_directory)
    config_path = agent_directory / "config.yaml"
    if not config_path.exists():
        raise FileNotFoundError(
            f"Configuration file not found at {config_path}. Please ensure the configuration has been set.")
    with open(config_path, "r") as f:
        agent_config = yaml.unsafe_load(f)
    agent = FQLAgent.create(agent_config["seed"], sample_batch["observations"], sample_batch["actions"], agent_config,
                            population=population, slot=slot)
    agent_path = agent_directory / (agent_filename + agent_extension)
    if not agent_path.exists():
        raise FileNotFoundError(f"Checkpoint not found at {agent_path}. Please ensure the agent has been trained.")
    with open(agent_path, "rb") as f:
        state = pickle.load(f)
    return from_state_dict(agent, state["agent"])
