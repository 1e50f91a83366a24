"""Helpers for chains of environment wrappers (pyrado/environment_wrappers/utils.py:29-200).  The real environment is
always at the end of the chain."""
from .base import EnvWrapper


def all_envs(env):
    """Iterate over the chain, outermost first."""
    yield env
    while isinstance(env, EnvWrapper):
        env = env.wrapped_env
        yield env


def typed_env(env, tp):
    """First element of the chain that is an instance of `tp`, or None."""
    return next((e for e in all_envs(env) if isinstance(e, tp)), None)


def attr_env(env, attr: str):
    """First element of the chain that has the attribute, or None."""
    return next((e for e in all_envs(env) if hasattr(e, attr)), None)


def attr_env_get(env, attr: str):
    """Value of the attribute of the first element of the chain that has it, or None."""
    for e in all_envs(env):
        try:
            return getattr(e, attr)
        except AttributeError:
            continue
    return None


def inner_env(env):
    """The innermost (non-wrapper) environment."""
    while isinstance(env, EnvWrapper):
        env = env.wrapped_env
    return env


def _rebuild(head, op):
    """Walk from the outside in; `op(env)` returns a replacement (stop), or None (go deeper).  Wrappers above a
    replaced element are copied, untouched parts are shared."""
    res = op(head)
    if res is not None:
        return res
    if isinstance(head, EnvWrapper):
        below = _rebuild(head.wrapped_env, op)
        if below is not head.wrapped_env:
            return head.copy(wrapped_env=below)
    return head


def remove_env(stack, key_type):
    """The chain without its first wrapper of type `key_type` (the original chain is not modified)."""
    assert issubclass(key_type, EnvWrapper)
    return _rebuild(stack, lambda e: e.wrapped_env if isinstance(e, key_type) else None)


def insert_env_before(stack, key_type, insert_type, *args, **kwargs):
    """The chain with `insert_type(env, *args, **kwargs)` put right around the first `key_type` element (around the
    innermost env for `key_type=None`)."""
    assert key_type is None or issubclass(key_type, EnvWrapper)
    assert issubclass(insert_type, EnvWrapper)

    def op(e):
        hit = (not isinstance(e, EnvWrapper)) if key_type is None else isinstance(e, key_type)
        return insert_type(e, *args, **kwargs) if hit else None

    return _rebuild(stack, op)
