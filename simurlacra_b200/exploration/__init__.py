"""Step-based exploration (pyrado/exploration): additive action noise, fused into the rollout kernel for sampling."""
from .noise import DiagNormalNoise, UniformNoise
from .stochastic_action import NormalActNoiseExplStrat, StochasticActionExplStrat, UniformActNoiseExplStrat

__all__ = ["DiagNormalNoise", "UniformNoise", "StochasticActionExplStrat", "NormalActNoiseExplStrat",
           "UniformActNoiseExplStrat"]
