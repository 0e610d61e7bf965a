"""Same module path as /root/reference/envmodel/termination_predictor.py."""
from envmodel.baseline import TerminationPredictor  # noqa: F401
