"""Environment wrappers (pyrado/environment_wrappers); ActNorm, ObsNorm and the Gaussian observation / action noise fuse into the kernels."""
from .action_delay import ActDelayWrapper  # noqa: F401
from .action_noise import GaussianActNoiseWrapper  # noqa: F401
from .action_normalization import ActNormWrapper, ObsActCatWrapper, dclamp  # noqa: F401
from .base import EnvWrapper, EnvWrapperAct, EnvWrapperObs  # noqa: F401
from .domain_randomization import (  # noqa: F401
    DomainRandWrapper,
    DomainRandWrapperBuffer,
    DomainRandWrapperLive,
    MetaDomainRandWrapper,
    remove_all_dr_wrappers,
)
from .downsampling import DownsamplingWrapper  # noqa: F401
from .observation_noise import GaussianObsNoiseWrapper  # noqa: F401
from .observation_normalization import ObsNormWrapper  # noqa: F401
from .observation_partial import ObsPartialWrapper  # noqa: F401
from .observation_velfilter import ObsVelFiltWrapper  # noqa: F401
from .state_augmentation import StateAugmentationWrapper  # noqa: F401
from .utils import all_envs, attr_env, attr_env_get, inner_env, insert_env_before, remove_env, typed_env  # noqa: F401
