"""
SimPyEnv (pyrado/environments/pysim/base.py:44-289) as a BATCH of `num_envs` identical simulations that live on the
GPU.  The reference signatures are kept; with `num_envs == 1` (the default) the return types are the reference's
(`reset -> np.ndarray`, `step -> (np.ndarray, float, bool, dict)`, `state` a float64 ndarray), with `num_envs > 1`
the same calls take and return CUDA tensors with a leading batch dimension.

All arithmetic runs in the CUDA kernels behind the C-ABI (`_calc_constants`, `_step_dynamics`, reward, done,
observe, init-state sampling); this class only owns device buffers and host-side descriptors (spaces, task).
"""
from abc import abstractmethod
from copy import deepcopy
from typing import Optional, Union

import numpy as np
import torch

import simurlacra_b200 as pyrado

from ... import _lib as L
from ... import engine as E
from ...spaces.base import Space
from ...spaces.box import BoxSpace
from ...spaces.singular import SingularStateSpace
from ...tasks.base import Task
from ..sim_base import SimEnv


class Fusion:
    """What the env wrappers around a SimPyEnv fuse into its kernels (see environment_wrappers/)."""

    def __init__(self):
        self.act_norm = False
        self.noise_std = None
        self.noise_mean = None
        self.obs_norm = None        # (lb, ub) of an ObsNormWrapper
        self.env_act_noise = None   # (position, std, mean) of a GaussianActNoiseWrapper
        self.wrap = None            # dict of the stateful step wrappers the persistent rollout kernel runs (engine.make_cfg)


class SimPyEnv(SimEnv):
    kind: int = None           # SRE_* env kind
    _param_names: tuple = ()   # domain parameter names in the order of the C-ABI parameter vector
    _pass_init_box = False     # True: the class's init space is not the kernel's built-in one -> passed as a box

    def __init__(self, dt: float, max_steps: int = pyrado.inf, task_args: Optional[dict] = None, *, num_envs: int = 1,
                 device: Union[str, torch.device, None] = None, dtype: Optional[torch.dtype] = None):
        """
        :param dt: simulation step size [s]
        :param max_steps: maximum number of simulation steps
        :param task_args: arguments for the task construction
        :param num_envs: number of environments in the batch (engine extension; 1 keeps the reference types)
        :param device: CUDA device of the batch (default: current device)
        :param dtype: torch.float64 (verification build, reference-like) or torch.float32 (production build);
                      default float64 for a single env, float32 for a batch
        """
        super().__init__(dt, max_steps)
        if not isinstance(num_envs, int) or num_envs < 1:
            raise pyrado.ValueErr(given=num_envs, ge_constraint="1 (int)")
        self.num_envs = num_envs
        self._device = torch.device(device) if device is not None else None
        self._dtype = dtype if dtype is not None else (torch.float64 if num_envs == 1 else torch.float32)
        self._dims = E.dims(self.kind)
        self._flags = 0
        self._wild_init = 0

        # Domain parameters: one dict for the API, one [N, dp] device tensor once they differ per env
        self._domain_param = self.get_nominal_domain_param()
        self._params = None
        self._state_space = self._obs_space = self._act_space = self._init_space = None
        self._custom_init_space = False
        self._create_spaces()

        if not (isinstance(task_args, dict) or task_args is None):
            raise pyrado.TypeErr(given=task_args, expected_type=dict)
        self._task = self._create_task(task_args=dict() if task_args is None else task_args)

        # Device state, allocated at the first reset
        self._state = self._hidden = self._curr_step_t = None
        self.env_offset = 0   # global index of env 0 (multi-GPU shards set this)
        self._episode = 0     # Philox episode counter, advanced by every reset
        self._curr_act = np.zeros(self.act_space.shape)

    # ------------------------------------------------------------------ spaces / task / properties
    @property
    def device(self) -> torch.device:
        if self._device is None:
            if not torch.cuda.is_available():
                raise L.EngineError("simurlacra_b200 needs a CUDA device (there is no CPU fallback)")
            self._device = torch.device("cuda", torch.cuda.current_device())
        return self._device

    @property
    def dtype(self) -> torch.dtype:
        return self._dtype

    @property
    def state_space(self) -> Space:
        return self._state_space

    @property
    def obs_space(self) -> Space:
        return self._obs_space

    @property
    def init_space(self) -> Space:
        return self._init_space

    @init_space.setter
    def init_space(self, space: Space):
        if not isinstance(space, Space):
            raise pyrado.TypeErr(given=space, expected_type=Space)
        self._init_space = space
        self._custom_init_space = True

    @property
    def act_space(self) -> Space:
        return self._act_space

    @property
    def task(self) -> Task:
        return self._task

    @abstractmethod
    def _create_spaces(self):
        """Host-side descriptors of the spaces for the current (env 0) domain parameters."""
        raise NotImplementedError

    @abstractmethod
    def _create_task(self, task_args: dict) -> Task:
        raise NotImplementedError

    # ------------------------------------------------------------------ domain parameters
    @property
    def domain_param(self) -> dict:
        """A deep copy like the reference's getter (base.py:109-110); per-env values come back as float64 arrays."""
        if self._params is None:
            return deepcopy(self._domain_param)
        p = self._params.double().cpu().numpy()
        return {k: (float(p[0, i]) if self.num_envs == 1 else p[:, i].copy()) for i, k in enumerate(self._param_names)}

    @domain_param.setter
    def domain_param(self, domain_param: dict):
        if not isinstance(domain_param, dict):
            raise pyrado.TypeErr(given=domain_param, expected_type=dict)
        per_env = False
        for k, v in domain_param.items():
            if k not in self._param_names:
                continue  # wrappers strip their own entries before; unknown keys are ignored like dict.update + attrs
            arr = np.asarray(v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v, dtype=np.float64)
            if arr.ndim >= 1 and arr.size == self.num_envs and self.num_envs > 1:
                per_env = True
            elif arr.size != 1:
                raise pyrado.ShapeErr(msg=f"Domain parameter {k} must be a scalar or have one value per env!")
        if per_env or self._params is not None:
            if self._params is None:
                base = np.array([self._domain_param[k] for k in self._param_names], dtype=np.float64)
                host = np.tile(base, (self.num_envs, 1))
            else:
                host = self._params.double().cpu().numpy()
            for k, v in domain_param.items():
                if k in self._param_names:
                    arr = np.asarray(v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v, dtype=np.float64)
                    host[:, self._param_names.index(k)] = arr.reshape(-1) if arr.size > 1 else float(arr)
            self._params = torch.as_tensor(host, dtype=self._dtype).to(self.device).contiguous()
            for i, k in enumerate(self._param_names):
                self._domain_param[k] = float(host[0, i])
        else:
            for k, v in domain_param.items():
                if k in self._param_names:
                    self._domain_param[k] = float(np.asarray(v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v))
        # _calc_constants happens inside the kernels; spaces and the task follow the new parameters (base.py:112-124)
        self._create_spaces()
        self._task.reset(env_spec=self.spec)

    def set_param_tensor(self, params: torch.Tensor):
        """Engine extension: adopt a [num_envs, dp] device tensor of domain parameters (e.g. from the DR kernel)."""
        if params.shape != (self.num_envs, self._dims["dp"]):
            raise pyrado.ShapeErr(given=params, expected_match=(self.num_envs, self._dims["dp"]))
        self._params = params.to(self._dtype).contiguous()

    def randomize_params(self, randomizer, episode: Optional[int] = None, seed: Optional[int] = None):
        """Engine extension: draw a fresh parameter set per env from `randomizer` with one launch of the Philox
        sampler (DomainRandomizer.randomize + domain_param setter for the whole batch).  Keyed (seed, env_offset + i,
        episode), the key the rollout kernel uses when it samples the same randomizer in-flight."""
        if self._params is None:
            base = torch.as_tensor(self.base_param_vector(), dtype=self._dtype)
            self._params = base.repeat(self.num_envs, 1).to(self.device).contiguous()
        episode = self._episode + 1 if episode is None else episode  # the reset that follows advances to it
        E.sample_params(self._make_cfg(), randomizer.table(self._param_names), self._params,
                        pyrado.device_seed() if seed is None else seed, self.env_offset, episode)
        first = self._params[0].double().cpu().numpy()
        for i, k in enumerate(self._param_names):
            self._domain_param[k] = float(first[i])
        self._create_spaces()
        self._task.reset(env_spec=self.spec)

    def base_param_vector(self) -> np.ndarray:
        return np.array([self._domain_param[k] for k in self._param_names], dtype=np.float64)

    # ------------------------------------------------------------------ C-ABI configuration
    def _make_cfg(self, fusion: Optional[Fusion] = None) -> L.EnvCfg:
        task = L.Task()
        self._task.fill(task)
        init_box = None
        if (self._custom_init_space or self._pass_init_box) and type(self._init_space) in (BoxSpace, SingularStateSpace):
            init_box = self._init_space.bounds
        f = fusion or Fusion()
        return E.make_cfg(self.kind, self._dt, self._max_steps, dtype=self._dtype,
                          simple_dynamics=bool(self._flags & L.F_SIMPLE_DYN), wild_init=self._wild_init, task=task,
                          base_params=self.base_param_vector(), act_norm=f.act_norm, noise_std=f.noise_std,
                          noise_mean=f.noise_mean, init_box=init_box, flags=self._flags & ~L.F_SIMPLE_DYN,
                          obs_norm=f.obs_norm, env_act_noise=f.env_act_noise, wrap=f.wrap)

    def _alloc(self):
        if self._state is None:
            n, d = self.num_envs, self._dims
            self._state = torch.zeros(n, d["ds"], dtype=self._dtype, device=self.device)
            self._hidden = torch.zeros(n, max(d["nh"], 1), dtype=self._dtype, device=self.device)
            self._curr_step_t = torch.zeros(n, dtype=torch.int32, device=self.device)

    def _to_dev(self, x, cols: int) -> torch.Tensor:
        if isinstance(x, torch.Tensor):
            t = x.to(device=self.device, dtype=self._dtype)
        else:
            t = torch.as_tensor(np.asarray(x, dtype=np.float64), dtype=self._dtype).to(self.device)
        return t.reshape(-1, cols).contiguous()

    def _out(self, t: torch.Tensor):
        """Reference types for a single env, device tensors for a batch."""
        if self.num_envs == 1:
            return t[0].double().cpu().numpy()
        return t

    # ------------------------------------------------------------------ state access
    @property
    def state(self):
        """`env.state` (public attribute in the reference): float64 ndarray for one env, device tensor for a batch."""
        if self._state is None:
            return None
        return self._state[0].double().cpu().numpy() if self.num_envs == 1 else self._state

    @state.setter
    def state(self, value):
        self._alloc()
        self._state.copy_(self._to_dev(value, self._dims["ds"]).expand_as(self._state))

    @property
    def hidden_state(self):
        """th_ddot carry (cart-pole) / plate angles (ball balancer): hidden attributes in the reference."""
        return None if self._hidden is None else self._hidden[:, : self._dims["nh"]]

    @property
    def curr_step(self):
        if self._curr_step_t is None or self.num_envs == 1:
            return self._curr_step
        return self._curr_step_t & L.STEP_COUNT_MASK  # bit 30 carries the FinalRewTask "yielded" flag (include/sre.h)

    def observe(self, state) -> np.ndarray:
        """Noise-free observation of a given state (Env.observe, environments/base.py:203 and the overrides)."""
        s = self._to_dev(state, self._dims["ds"])
        obs = E.observe(self._make_cfg(), s)
        if not isinstance(state, torch.Tensor):
            return obs.double().cpu().numpy().reshape(np.shape(state)[:-1] + (self._dims["do"],))
        return obs

    # ------------------------------------------------------------------ reset / step (base.py:166-241)
    def reset(self, init_state=None, domain_param: dict = None):
        return self._reset_impl(init_state, domain_param, None)

    def _reset_impl(self, init_state, domain_param, fusion: Optional[Fusion]):
        self._curr_step = 0
        if domain_param is not None:
            self.domain_param = domain_param
        self._alloc()
        cfg = self._make_cfg(fusion)
        seed = pyrado.device_seed()
        self._episode += 1
        if init_state is None:
            if self._custom_init_space and type(self._init_space) is not BoxSpace:
                # exotic user-provided init space: host sampling like the reference (init_space.sample_uniform())
                init = self._to_dev(np.stack([self._init_space.sample_uniform() for _ in range(self.num_envs)]),
                                    self._init_space.flat_dim)
            else:
                init = E.sample_init(cfg, self._params, self.num_envs, seed, self.env_offset, self._episode,
                                     device=self.device)
        else:
            if not isinstance(init_state, (np.ndarray, torch.Tensor)):
                try:
                    init_state = np.asarray(init_state, dtype=np.float64)
                except Exception:
                    raise pyrado.TypeErr(given=init_state, expected_type=np.ndarray)
            last = init_state.shape[-1] if init_state.ndim else 1
            if last not in (self._dims["ds"], self._dims["dinit"]):
                raise pyrado.ShapeErr(given=init_state, expected_match=self.state_space)
            init = self._to_dev(init_state, last)
            if init.shape[0] == 1 and self.num_envs > 1:
                init = init.expand(self.num_envs, -1).contiguous()
            if last != self._dims["ds"] and self.num_envs == 1:
                if not self.init_space.contains(init[0].double().cpu().numpy(), verbose=True):
                    print("The init state is not within init state space.")  # warning only, base.py:191-192
        obs = torch.empty(self.num_envs, self._dims["do"], dtype=self._dtype, device=self.device)
        E.reset(cfg, self._params, init, self._state, self._hidden, self._curr_step_t, obs, None, seed, self.env_offset,
                self._episode)
        self._task.reset(env_spec=self.spec)
        return self._out(obs)

    def step(self, act) -> tuple:
        return self._step_impl(act, None)

    def _reset_chain(self, init_state, domain_param, fusion):  # entry points of the wrapper chain
        return self._reset_impl(init_state, domain_param, fusion)

    def _step_chain(self, act, fusion) -> tuple:
        return self._step_impl(act, fusion)

    def _step_impl(self, act, fusion: Optional[Fusion]) -> tuple:
        if self._state is None:
            raise RuntimeError("reset() must be called before step()")
        cfg = self._make_cfg(fusion)
        n, d = self.num_envs, self._dims
        a = self._to_dev(act, d["da"])
        if a.shape[0] != n:
            raise pyrado.ShapeErr(given=act, expected_match=self.act_space)
        obs = torch.empty(n, d["do"], dtype=self._dtype, device=self.device)
        rew = torch.empty(n, dtype=self._dtype, device=self.device)
        done = torch.empty(n, dtype=torch.uint8, device=self.device)
        status = torch.empty(n, dtype=torch.int32, device=self.device)
        E.step(cfg, self._params, a, self._state, self._hidden, self._curr_step_t, obs, rew, done, status,
               pyrado.device_seed(), self.env_offset, self._episode)
        self._curr_step += 1
        if n == 1:
            st = int(status[0])
            if st & (L.ST_NAN_STATE | L.ST_NAN_ACT):  # the reference raises in BoxSpace.contains (spaces/box.py:142)
                raise pyrado.ValueErr(msg=f"At least one value is NaN! (status {st})")
            self._curr_rew = float(rew[0])
            self._curr_act = np.asarray(act, dtype=np.float64).reshape(-1)
            return obs[0].double().cpu().numpy(), self._curr_rew, bool(done[0]), dict()
        return obs, rew, done.bool(), dict(status=status)

    def _state_from_init(self, init_state: np.ndarray):
        assert self._init_space.shape == self._state_space.shape
        return init_state
