import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "flow-q-learning_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


# the tests train on generated data of OGBench shape (no datasets exist on these boxes): explicit opt-in
os.environ.setdefault("FQLPOP_SYNTHETIC", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
