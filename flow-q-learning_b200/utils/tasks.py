"""Task-name helpers (/root/reference/utils/tasks.py:1-15)."""


def _family(env_name: str):
    for name in ("antsoccer", "cube"):
        if env_name.startswith(name):
            return name
    return None


def get_task_title(env_name: str) -> str:
    fam = _family(env_name)
    return fam.capitalize() if fam else "Unknown Task"


def get_task_filename(env_name: str) -> str:
    return _family(env_name) or "unknown"
