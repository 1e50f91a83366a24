"""
Policy base class (pyrado/policies/base.py:57-175).  A policy is a torch ``nn.Module`` for the learner (autograd,
optimisers, state_dict) exactly as in the reference; for sampling it additionally compiles itself into the C-ABI
policy descriptor (``engine_desc``) so that the rollout kernel evaluates it fused with the simulation step, and
``forward_engine`` evaluates a batch of observations with the stand-alone CUDA kernel (``sre_policy_forward``).
"""
from abc import ABC, abstractmethod
from typing import Optional
from warnings import warn

import torch
import torch.nn as nn
from torch.nn.utils import convert_parameters as cp

import simurlacra_b200 as pyrado

from .. import _lib as L
from .. import engine as E
from ..utils.data_types import EnvSpec


def _get_or_create_grad(t):
    if t.grad is None:
        t.grad = torch.zeros_like(t)
    return t.grad


class Policy(nn.Module, ABC):
    """Base class for all policies"""

    name: str = None  # unique identifier

    def __init__(self, spec: EnvSpec, use_cuda: bool = False):
        if not isinstance(spec, EnvSpec):
            raise pyrado.TypeErr(given=spec, expected_type=EnvSpec)
        if not isinstance(use_cuda, bool):
            raise pyrado.TypeErr(given=use_cuda, expected_type=bool)
        super().__init__()
        self._env_spec = spec
        if use_cuda and not torch.cuda.is_available():
            warn("Tried to run on CUDA, but it is not available. Falling back to CPU.")
        self._device = "cuda" if use_cuda and torch.cuda.is_available() else "cpu"
        # arithmetic of the fused hidden layer (engine extension): L.PREC_TF32X3 (default: tcgen05 tensor cores with the
        # 3 x TF32 split, fp32-level accuracy; applies to float32 [in, 64, 64, out] tanh networks, every other shape and
        # the float64 build run the CUDA-core FMA path whatever this says), L.PREC_FP32 (CUDA-core FMA) or L.PREC_TF32
        self.engine_precision = L.PREC_TF32X3
        self._pinned = None
        self._pinned_free = None  # CUDA event: the last H2D copy out of the pinned staging buffer has completed
        self._dev_weights = {}

    @property
    def device(self) -> str:
        return self._device

    @property
    def env_spec(self) -> EnvSpec:
        return self._env_spec

    @property
    def param_values(self) -> torch.Tensor:
        """The parameters as one flat vector (a copy); assigning writes them back."""
        return cp.parameters_to_vector(self.parameters())

    @param_values.setter
    def param_values(self, param: torch.Tensor):
        if not self.param_values.shape == param.shape:
            raise pyrado.ShapeErr(given=param, expected_match=self.param_values)
        cp.vector_to_parameters(param, self.parameters())

    @property
    def param_grad(self) -> torch.Tensor:
        return cp.parameters_to_vector(_get_or_create_grad(p) for p in self.parameters())

    @param_grad.setter
    def param_grad(self, param):
        cp.vector_to_parameters(param, (_get_or_create_grad(p) for p in self.parameters()))

    @property
    def num_param(self) -> int:
        return sum(p.data.numel() for p in self.parameters())

    @property
    def is_recurrent(self) -> bool:
        return False

    def init_hidden(self, batch_size: Optional[int] = None) -> torch.Tensor:
        raise AttributeError("Only recurrent policies should use the init_hidden() method.")

    @abstractmethod
    def init_param(self, init_values: Optional[torch.Tensor] = None, **kwargs):
        raise NotImplementedError

    def reset(self, **kwargs):
        """Called at the start of a rollout; nothing to do for feed-forward policies."""

    @abstractmethod
    def forward(self, *args, **kwargs) -> torch.Tensor:
        raise NotImplementedError

    def evaluate(self, rollout, hidden_states_name: str = "hidden_states") -> torch.Tensor:
        """Re-evaluate a rollout and return a derivable action tensor (base.py:177-190)."""
        self.eval()
        res = self(rollout.get_data_values("observations", truncate_last=True))
        self.train()
        return res

    # ------------------------------------------------------------------ engine side
    def engine_desc(self, device, param_sets=None, group_size: int = 0) -> Optional[L.Policy]:
        """This policy as a C-ABI ``sre_policy`` (weights copied to `device` as one float32 vector), or None if the
        kernels do not implement it; samplers then evaluate the module between step-kernel launches.
        With `param_sets` (float32 CUDA [n_sets, num_param]) and `group_size`, lane n of a rollout launch evaluates
        parameter set n // group_size instead of the module's own parameters (parameter-exploration sampling)."""
        return None

    def _engine_weights(self, device) -> torch.Tensor:
        """The flat float32 parameter vector on `device`.  Host-resident parameters travel through a pinned staging
        buffer into a persistent device buffer (one asynchronous H2D copy per call = the per-`sample()` policy
        update that the reference does by pickling the state_dict to every worker)."""
        with torch.no_grad():
            v = self.param_values.detach()
            self._host_weights = None
            if v.is_cuda:
                return v.to(device=device, dtype=torch.float32).contiguous()
            device = torch.device(device)
            if self._pinned is None or self._pinned.numel() != v.numel():
                self._pinned = torch.empty(v.numel(), dtype=torch.float32).pin_memory()
                self._pinned_free = None
            if self._pinned_free is not None:
                # an earlier asynchronous copy may still be queued behind a rollout launch: the staging buffer must not
                # change under it (the earlier launch would silently run with these later weights)
                self._pinned_free.synchronize()
            self._pinned.copy_(v)
            self._host_weights = self._pinned  # the same parameters on the host: sre_policy.weights_host (read at launch time)
            dev = self._dev_weights.get(device)
            if dev is None or dev.numel() != v.numel():
                dev = self._dev_weights[device] = torch.empty(v.numel(), dtype=torch.float32, device=device)
            dev.copy_(self._pinned, non_blocking=True)
            self._pinned_free = torch.cuda.Event()
            self._pinned_free.record(torch.cuda.current_stream(device))
            return dev

    def forward_engine(self, obs: torch.Tensor) -> torch.Tensor:
        """Batch forward pass with the CUDA kernel: obs float32 CUDA [N, obs_dim] -> act float32 [N, act_dim]."""
        desc = self.engine_desc(obs.device)
        if desc is None:
            raise L.EngineError(f"{type(self).__name__} has no CUDA kernel")
        return E.policy_forward(desc, obs.to(torch.float32).contiguous(), self._env_spec.act_space.flat_dim)
