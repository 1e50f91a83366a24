"""
Default randomizers of the pysim environments (pyrado/domain_randomization/default_randomizers.py:46-456), written as
tables: one row per parameter = (name, distribution, spread as a fraction of the nominal value or absolute, clip_lo,
clip_up).  ``N`` rows are ``NormalDomainParam(mean=nom, std=nom/div)``, ``U`` rows ``UniformDomainParam(mean=nom,
halfspan=|nom|/div)``, ``Ua`` rows uniform with an absolute half span around the nominal value.
"""
from typing import Callable, Dict, Tuple, Union

import numpy as np

import simurlacra_b200 as pyrado

from ..environment_wrappers.base import EnvWrapper
from ..environment_wrappers.utils import inner_env
from ..environments.sim_base import SimEnv
from .domain_parameter import NormalDomainParam, UniformDomainParam
from .domain_randomizer import DomainRandomizer

DEFAULT_RANDOMIZER_REGISTRY: Dict[Tuple[str, str], Callable[[], DomainRandomizer]] = {}
_INF = pyrado.inf


def default_randomizer(env_module: str, env_class: str):
    """Decorator registering a default-randomizer provider for an env type referenced by name (:46-69)."""

    def register(func):
        DEFAULT_RANDOMIZER_REGISTRY[(env_module, env_class)] = func
        return func

    return register


def _from_rows(dp_nom: dict, rows) -> DomainRandomizer:
    params = []
    for name, kind, spread, lo, up in rows:
        nom = dp_nom[name]
        kw = dict(name=name, clip_lo=lo, clip_up=up)
        if kind == "N":
            params.append(NormalDomainParam(mean=nom, std=nom / spread, **kw))
        elif kind == "U":
            params.append(UniformDomainParam(mean=nom, halfspan=abs(nom) / spread, **kw))
        else:
            params.append(UniformDomainParam(mean=nom, halfspan=spread, **kw))
    return DomainRandomizer(*params)


def create_default_randomizer(env: Union[SimEnv, EnvWrapper]) -> DomainRandomizer:
    """The default randomizer for the innermost env's type, searched along its MRO (:72-89)."""
    env_type = type(inner_env(env))
    for cand in env_type.__mro__:
        fcn = DEFAULT_RANDOMIZER_REGISTRY.get((cand.__module__, cand.__name__))
        if fcn:
            return fcn()
    raise pyrado.ValueErr(msg=f"No default randomizer settings for env of type {env_type}!")


def create_conservative_randomizer(env: Union[SimEnv, EnvWrapper]) -> DomainRandomizer:
    """Half the spread of the default randomizer (:92-103)."""
    randomizer = create_default_randomizer(env)
    randomizer.rescale_distr_param("std", 0.5)
    randomizer.rescale_distr_param("halfspan", 0.5)
    return randomizer


def create_zero_var_randomizer(env: Union[SimEnv, EnvWrapper], eps: float = 1e-8) -> DomainRandomizer:
    """(Almost) always the nominal values; the variance is eps-scaled, not zero (:106-122)."""
    randomizer = create_default_randomizer(env)
    randomizer.rescale_distr_param("std", np.sqrt(eps))
    randomizer.rescale_distr_param("halfspan", np.sqrt(eps))
    return randomizer


def create_empty_randomizer() -> DomainRandomizer:
    return DomainRandomizer()


_PYSIM = "simurlacra_b200.environments.pysim."


@default_randomizer(_PYSIM + "ball_on_beam", "BallOnBeamSim")
def create_default_randomizer_bob() -> DomainRandomizer:  # :164-188
    from ..environments.pysim.ball_on_beam import BallOnBeamSim

    nom = dict(BallOnBeamSim.get_nominal_domain_param(), ang_offset=0.0)  # the angle offset is centred on zero
    rows = [("gravity_const", "N", 10, 1e-4, _INF), ("ball_mass", "N", 5, 1e-4, _INF), ("ball_radius", "N", 5, 1e-4, _INF),
            ("beam_mass", "N", 5, 1e-3, _INF), ("beam_length", "N", 5, 1e-3, _INF), ("beam_thickness", "N", 5, 1e-3, _INF),
            ("friction_coeff", "U", 1, 0, _INF), ("ang_offset", "Ua", 0.1 / 180 * np.pi, -_INF, _INF)]
    return _from_rows(nom, rows)


@default_randomizer(_PYSIM + "one_mass_oscillator", "OneMassOscillatorSim")
def create_default_randomizer_omo() -> DomainRandomizer:  # :191-204
    from ..environments.pysim.one_mass_oscillator import OneMassOscillatorSim

    nom = OneMassOscillatorSim.get_nominal_domain_param()
    return _from_rows(nom, [(k, "N", 3, 1e-3, _INF) for k in ("mass", "stiffness", "damping")])


@default_randomizer(_PYSIM + "pendulum", "PendulumSim")
def create_default_randomizer_pend() -> DomainRandomizer:  # :208-230
    from ..environments.pysim.pendulum import PendulumSim

    nom = PendulumSim.get_nominal_domain_param()
    return _from_rows(nom, [(k, "N", 10, 1e-3, _INF) for k in ("gravity_const", "pole_mass", "pole_length", "pole_damping",
                                                              "torque_thold")])


@default_randomizer(_PYSIM + "quanser_ball_balancer", "QBallBalancerSim")
def create_default_randomizer_qbb() -> DomainRandomizer:  # :233-307
    from ..environments.pysim.quanser_ball_balancer import QBallBalancerSim

    nom = QBallBalancerSim.get_nominal_domain_param()
    six_deg = 6.0 / 180 * np.pi
    rows = [("gravity_const", "N", 10, 1e-4, _INF), ("ball_mass", "N", 5, 1e-4, _INF), ("ball_radius", "N", 5, 1e-3, _INF),
            ("plate_length", "N", 5, 5e-2, _INF), ("arm_radius", "N", 5, 1e-4, _INF), ("gear_ratio", "N", 4, 1e-2, _INF),
            ("load_inertia", "N", 4, 1e-6, _INF), ("motor_inertia", "N", 4, 1e-9, _INF),
            ("motor_back_emf", "N", 4, 1e-4, _INF), ("motor_resistance", "N", 4, 1e-4, _INF),
            ("gear_efficiency", "U", 4, 1e-4, 1), ("motor_efficiency", "U", 4, 1e-4, 1),
            ("combined_damping", "U", 4, 1e-4, _INF), ("ball_damping", "U", 4, 1e-4, _INF),
            ("voltage_thold_x_pos", "U", 3, -_INF, _INF), ("voltage_thold_x_neg", "U", 3, -_INF, _INF),
            ("voltage_thold_y_pos", "U", 3, -_INF, _INF), ("voltage_thold_y_neg", "U", 3, -_INF, _INF),
            ("offset_th_x", "Ua", six_deg, -_INF, _INF), ("offset_th_y", "Ua", six_deg, -_INF, _INF)]
    return _from_rows(nom, rows)


@default_randomizer(_PYSIM + "quanser_cartpole", "QCartPoleStabSim")
@default_randomizer(_PYSIM + "quanser_cartpole", "QCartPoleSwingUpSim")
def create_default_randomizer_qcp() -> DomainRandomizer:  # :310-371 (short-pole nominal values)
    from ..environments.pysim.quanser_cartpole import QCartPoleSwingUpSim

    nom = QCartPoleSwingUpSim.get_nominal_domain_param(long=False)
    rows = [("gravity_const", "N", 10, 1e-4, _INF), ("cart_mass", "N", 5, 1e-4, _INF), ("pole_mass", "N", 5, 1e-4, _INF),
            ("rail_length", "N", 5, 1e-2, _INF), ("pole_length", "N", 5, 1e-2, _INF),
            ("motor_efficiency", "U", 4, 1e-4, 1), ("gear_efficiency", "U", 4, 1e-4, 1),
            ("gear_ratio", "N", 4, 1e-4, _INF), ("motor_inertia", "N", 4, 1e-9, _INF),
            ("pinion_radius", "N", 5, 1e-4, _INF), ("motor_resistance", "N", 4, 1e-4, _INF),
            ("motor_back_emf", "N", 4, 1e-4, _INF), ("combined_damping", "U", 4, 1e-4, _INF),
            ("pole_damping", "U", 4, 1e-4, _INF), ("cart_friction_coeff", "U", 2, 0, _INF)]
    return _from_rows(nom, rows)


@default_randomizer(_PYSIM + "quanser_qube", "QQubeSwingUpSim")
@default_randomizer(_PYSIM + "quanser_qube", "QQubeStabSim")
def create_default_randomizer_qq() -> DomainRandomizer:  # :374-416
    from ..environments.pysim.quanser_qube import QQubeSwingUpSim

    nom = QQubeSwingUpSim.get_nominal_domain_param()
    rows = [("gravity_const", "N", 10, 1e-3, _INF), ("motor_resistance", "N", 5, 1e-3, _INF),
            ("motor_back_emf", "N", 5, 1e-4, _INF), ("mass_rot_pole", "N", 5, 1e-4, _INF),
            ("length_rot_pole", "N", 5, 1e-4, _INF), ("damping_rot_pole", "N", 4, 1e-9, _INF),
            ("mass_pend_pole", "N", 5, 1e-4, _INF), ("length_pend_pole", "N", 5, 1e-4, _INF),
            ("damping_pend_pole", "N", 4, 1e-9, _INF)]
    return _from_rows(nom, rows)


def create_uniform_masses_lengths_randomizer_qq(frac_halfspan: float) -> DomainRandomizer:
    """Uniform randomizer over the Qube's masses and lengths, half span = nominal / frac_halfspan (:419-456)."""
    from ..environments.pysim.quanser_qube import QQubeSwingUpSim

    nom = QQubeSwingUpSim.get_nominal_domain_param()
    rows = [("mass_pend_pole", "U", frac_halfspan, 1e-3, _INF), ("mass_rot_pole", "U", frac_halfspan, 1e-3, _INF),
            ("length_rot_pole", "U", frac_halfspan, 1e-2, _INF), ("length_pend_pole", "U", frac_halfspan, 1e-2, _INF)]
    return _from_rows(nom, rows)
