"""Sampling (pyrado/sampling): StepSequence, rollout(), ParallelRolloutSampler, ParameterExplorationSampler, the differentiable cart-pole path."""
from .parallel_rollout_sampler import ParallelRolloutSampler, ParallelRolloutSamplerTensor  # noqa: F401
from .parameter_exploration_sampler import ParameterExplorationSampler, ParameterSample, ParameterSamplingResult  # noqa: F401
from .rollout import rollout  # noqa: F401
from .rollout_tensor import rollout_tensor_full  # noqa: F401
from .step_sequence import StepSequence, discounted_reverse_cumsum, discounted_value, discounted_values, gae_returns  # noqa: F401
from . import postprocess  # noqa: F401
