"""tasks/utils.py:32-34: the success function used by all five pysim envs."""


def never_succeeded(*args, **kwargs) -> bool:
    return False
