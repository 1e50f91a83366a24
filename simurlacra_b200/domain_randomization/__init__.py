"""Domain parameters and randomizers (pyrado/domain_randomization)."""
from .domain_parameter import BernoulliDomainParam, DomainParam, NormalDomainParam, UniformDomainParam  # noqa: F401
from .domain_randomizer import DomainRandomizer  # noqa: F401
