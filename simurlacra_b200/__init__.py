"""
simurlacra_b200 -- B200-native batched rollout engine for Pyrado's pysim environments (SimuRLacra).

The package mirrors the reference's API for this one path (``SimPyEnv.reset/step``, domain parameters,
``DomainRandomizer`` / ``DomainRandWrapper``, ``rollout()``, ``ParallelRolloutSampler``, the cart-pole
``step_diff_state``) on top of hand-written CUDA kernels for sm_100a behind a C-ABI (``include/sre.h``).
There is no CPU fallback.
"""
__version__ = "0.1.0"
