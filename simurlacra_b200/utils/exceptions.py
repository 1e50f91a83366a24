"""Exception types with the reference's names and keyword interface (pyrado/utils/exceptions.py:86,130,197,275)."""


def _fmt(**kw):
    return ", ".join(f"{k}={v!r}" for k, v in kw.items() if v is not None)


class TypeErr(TypeError):
    def __init__(self, *, given=None, expected_type=None, msg: str = None):
        super().__init__(msg or f"Expected type {expected_type}, but received {type(given)}!")


class ValueErr(ValueError):
    def __init__(self, *, given=None, given_name=None, eq_constraint=None, l_constraint=None, le_constraint=None,
                 g_constraint=None, ge_constraint=None, msg: str = None):
        if msg is None:
            cons = _fmt(eq=eq_constraint, l=l_constraint, le=le_constraint, g=g_constraint, ge=ge_constraint)
            msg = f"The value {given_name or ''} = {given} violates the constraint(s): {cons}!"
        super().__init__(msg)


class ShapeErr(RuntimeError):
    def __init__(self, *, given=None, given_name=None, expected_match=None, msg: str = None):
        if msg is None:
            gs = getattr(given, "shape", given)
            es = getattr(expected_match, "shape", expected_match)
            msg = f"The shape {gs} of {given_name or 'the given object'} does not match the expected shape {es}!"
        super().__init__(msg)


class KeyErr(KeyError):
    def __init__(self, *, keys=None, container=None, msg: str = None):
        super().__init__(msg or f"The key(s) {keys} was/were not found in the container {type(container)}!")


class PathErr(FileNotFoundError):
    def __init__(self, *, given=None, msg: str = None):
        super().__init__(msg or f"The path {given} is not a valid file or directory!")
