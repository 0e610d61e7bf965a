"""Strategy that never prunes (/root/reference/hpo/identity.py:7-15)."""
from hpo.strategy import HpoStrategy


class Identity(HpoStrategy):
    def update(self, candidate, performance):
        return None

    def sample(self):
        return self.population
