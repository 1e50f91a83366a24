"""
simurlacra_b200 -- B200-native batched rollout engine for Pyrado's pysim environments (SimuRLacra).

The package mirrors the reference's API for this one path (``SimPyEnv.reset/step``, domain parameters,
``DomainRandomizer`` / ``DomainRandWrapper``, ``rollout()``, ``ParallelRolloutSampler``, the cart-pole
``step_diff_state``) on top of hand-written CUDA kernels for sm_100a behind a C-ABI (``include/sre.h``).
There is no CPU fallback.  Top-level names follow ``pyrado/__init__.py``.
"""
import hashlib
import random
from typing import Optional

import numpy as np

from .utils.exceptions import KeyErr, PathErr, ShapeErr, TypeErr, ValueErr  # noqa: F401

__version__ = "0.1.0"

inf = float("inf")  # pyrado.inf
nan = float("nan")
NO_SEED = None

_BASE_SEED = None
_DEVICE_SEED = 0  # key of the device-side Philox streams: the last hashed seed set with set_seed()


def device_seed() -> int:
    return _DEVICE_SEED


def set_device_seed(seed: int):
    """Key of the device-side Philox streams (engine extension; `set_seed` sets it to the hashed seed)."""
    global _DEVICE_SEED
    _DEVICE_SEED = int(seed) & 0xFFFFFFFFFFFFFFFF


def seed_hash(seed: int, sub_seed: Optional[int] = None, sub_sub_seed: Optional[int] = None) -> int:
    """md5("base-sub-subsub") mod 2^32 (pyrado/__init__.py:168); pinned by Pyrado/tests/test_set_seed.py:34-55."""
    sub_seed = 0 if sub_seed is None else sub_seed
    sub_sub_seed = 0 if sub_sub_seed is None else sub_sub_seed
    return int(hashlib.md5(f"{seed}-{sub_seed}-{sub_sub_seed}".encode()).hexdigest(), 16) % (2 ** 32)


def set_seed(seed: Optional[int], sub_seed: int = None, sub_sub_seed: int = None, verbose: bool = False) -> Optional[int]:
    """
    Same contract as ``pyrado.set_seed`` (pyrado/__init__.py:135-186): hashes the triple into a 32-bit seed and seeds
    the host generators (random, numpy, torch).  The device-side Philox streams are keyed separately per call by
    (seed, global env index, episode), see ``sampling``.
    """
    global _BASE_SEED, _DEVICE_SEED
    if not isinstance(seed, int):
        if verbose:
            print(f"Base seed {seed} is not an integer -- the random number generators' seeds were not set.")
        return None
    s = seed_hash(seed, sub_seed, sub_sub_seed)
    random.seed(s)
    np.random.seed(s)
    import torch

    torch.manual_seed(s)
    _BASE_SEED = seed
    _DEVICE_SEED = s
    if verbose:
        print(f"Set the random number generators' seed to {s} (base seed {seed}, sub-seed {sub_seed or 0}, "
              f"sub-sub-seed {sub_sub_seed or 0}).")
    return s


def get_base_seed() -> Optional[int]:
    return _BASE_SEED
