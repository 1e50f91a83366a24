"""
TEST INFRASTRUCTURE ONLY -- import shims that let the *unmodified* reference
(`/root/reference/Pyrado/pyrado`) be imported in this container so it can be run as the
parity oracle and to generate the golden vectors under `tests/golden/`.

Nothing in the product package (`simurlacra_b200/`) may import this module.  It is only used by
`oracle/gen_golden.py` (run once, here, where `/root/reference` exists) -- the GPU box has no
`/root/reference`, so tests and the bench never call this at run time.

What is shimmed:
  * absent third-party modules: init_args_serializer, colorama, ipdb, matplotlib*, seaborn, sbi*,
    optuna, prettyprinter, dtw, gym*, tensorboardX, pandas-free paths stay untouched
  * `np.float` (removed alias used at quanser_cartpole.py:164)
  * `torch.solve` (removed in torch >= 1.13, used at quanser_cartpole.py:422)
"""
import importlib.abc
import importlib.machinery
import sys
import types

REF_ROOT = "/root/reference/Pyrado"

_STUB_ROOTS = (
    "matplotlib", "seaborn", "sbi", "optuna", "prettyprinter", "dtw", "gym", "tensorboardX",
    "mpl_toolkits", "pandas_profiling", "tikzplotlib", "scikits", "vortex", "rcsenv", "mujoco_py", "pytest_lazyfixture",
)


class _Dummy:
    """Permissive stand-in class: subclassable, callable, attribute access returns itself."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Dummy()

    def __getattr__(self, item):
        if item.startswith("__"):
            raise AttributeError(item)
        return _Dummy()

    def __iter__(self):
        return iter(())

    def __mro_entries__(self, bases):
        return (_Dummy,)


class _StubModule(types.ModuleType):
    def __getattr__(self, item):
        if item.startswith("__"):
            raise AttributeError(item)
        if item[:1].isupper():
            cls = type(item, (_Dummy,), {})
            setattr(self, item, cls)
            return cls
        full = self.__name__ + "." + item
        if full in sys.modules:
            return sys.modules[full]
        # functions / submodules: return a permissive callable stub module
        mod = _CallableStubModule(full)
        mod.__path__ = []
        sys.modules[full] = mod
        setattr(self, item, mod)
        return mod


class _CallableStubModule(_StubModule):
    def __call__(self, *a, **k):
        return _Dummy()


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path, target=None):
        root = fullname.split(".")[0]
        if root in _STUB_ROOTS:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        mod = _CallableStubModule(spec.name)
        mod.__path__ = []
        return mod

    def exec_module(self, module):
        pass


def _make_serializable():
    mod = types.ModuleType("init_args_serializer")
    sub = types.ModuleType("init_args_serializer.serializable")

    class Serializable:
        """Minimal re-statement of the un-vendored `init_args_serializer.Serializable` (no arithmetic)."""

        def _init(self, locals_):
            if getattr(self, "_serializable_initialized", False):
                return
            import inspect

            spec = inspect.getfullargspec(self.__init__)
            args = [locals_[a] for a in spec.args[1:] if a in locals_]
            kwargs = dict(locals_.get(spec.varkw, {})) if spec.varkw else {}
            if spec.varargs:
                args += list(locals_.get(spec.varargs, ()))
            for k in spec.kwonlyargs:
                if k in locals_:
                    kwargs[k] = locals_[k]
            self._ser_args = tuple(args)
            self._ser_kwargs = kwargs
            self._serializable_initialized = True

        def _get_state(self, state_dict):
            pass

        def _set_state(self, state_dict, copying=False):
            pass

        def __getstate__(self):
            d = {"__args": self._ser_args, "__kwargs": self._ser_kwargs}
            sd = {}
            self._get_state(sd)
            d["state"] = sd
            return d

        def __setstate__(self, d):
            obj = type(self)(*d["__args"], **d["__kwargs"])
            self.__dict__.update(obj.__dict__)
            self._set_state(d.get("state", {}))

        def copy(self, **overrides):
            import inspect

            spec = inspect.getfullargspec(self.__init__)
            kw = dict(zip(spec.args[1:], self._ser_args))
            kw.update(self._ser_kwargs)
            kw.update(overrides)
            return type(self)(**kw)

    mod.Serializable = Serializable
    sub.Serializable = Serializable
    mod.serializable = sub
    sys.modules["init_args_serializer"] = mod
    sys.modules["init_args_serializer.serializable"] = sub


def _make_colorama():
    mod = types.ModuleType("colorama")

    class _Codes:
        def __getattr__(self, item):
            return ""

    mod.init = lambda *a, **k: None
    mod.Style = _Codes()
    mod.Fore = _Codes()
    mod.Back = _Codes()
    sys.modules["colorama"] = mod


def _make_ipdb():
    mod = types.ModuleType("ipdb")

    def set_trace(*a, **k):
        raise RuntimeError("ipdb.set_trace() reached inside the reference")

    mod.set_trace = set_trace
    sys.modules["ipdb"] = mod


_installed = False


def install():
    """Install all shims and put the reference on sys.path. Idempotent."""
    global _installed
    if _installed:
        return
    import numpy as np
    import torch

    _make_serializable()
    _make_colorama()
    _make_ipdb()
    sys.meta_path.insert(0, _StubFinder())
    if not hasattr(np, "float"):
        np.float = float
    # torch >= 1.13 keeps a `torch.solve` stub that only raises: always replace it
    torch.solve = lambda B, A: (torch.linalg.solve(A, B), None)
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    _installed = True
