"""
FeatureStack, the element-wise feature functions and the parametrised features MultFeat, ATan2Feat, RFFeat, RBFFeat
(pyrado/policies/features.py:42-445).  All of them are implemented by the Linear-policy kernel
(csrc/sre_policy.cuh ``linear_accumulate``); a stack holding any other callable still works as a torch module but is
evaluated outside the fused kernel.
"""
import functools
from typing import Callable, Optional, Sequence, Tuple, Union

import numpy as np
import torch

import simurlacra_b200 as pyrado


def const_feat(inp: torch.Tensor):
    if inp.ndim == 2:
        return torch.ones(inp.shape[0], 1).type_as(inp)
    return torch.tensor([1.0]).type_as(inp)


def identity_feat(inp: torch.Tensor):
    return inp.clone()


def sign_feat(inp: torch.Tensor):
    return torch.sign(inp)


def abs_feat(inp: torch.Tensor):
    return torch.abs(inp)


def squared_feat(inp: torch.Tensor):
    return torch.pow(inp, 2)


def cubic_feat(inp: torch.Tensor):
    return torch.pow(inp, 3)


def sig_feat(inp: torch.Tensor, scale: float = 1.0):
    return torch.sigmoid(scale * inp)


def bell_feat(inp: torch.Tensor, scale: float = 1.0):
    return torch.exp(-scale * torch.pow(inp, 2))


def sin_feat(inp: torch.Tensor):
    return torch.sin(inp)


def cos_feat(inp: torch.Tensor):
    return torch.cos(inp)


def sinsin_feat(inp: torch.Tensor):
    return torch.sin(inp) * torch.sin(inp)


def sincos_feat(inp: torch.Tensor):
    return torch.sin(inp) * torch.cos(inp)


class MultFeat:
    """Product of selected input dimensions (features.py:181-210): one feature."""

    def __init__(self, idcs: Tuple):
        if not isinstance(idcs, tuple):
            raise pyrado.TypeErr(given=idcs, expected_type=tuple)
        if not len(idcs) >= 2:
            raise pyrado.ShapeErr(msg="Provide at least provide two indices.")
        self._idcs = idcs

    def __str__(self):
        return "MultFeat (indices " + " ".join(str(i) for i in self._idcs) + ")"

    def __call__(self, inp: torch.Tensor) -> torch.Tensor:
        out = inp[..., self._idcs[0]]
        for i in self._idcs[1:]:
            out = out * inp[..., i]
        return out.unsqueeze(-1)

    def engine_code(self, n_obs: int):
        return ("mult", [len(self._idcs), *self._idcs])


class ATan2Feat:
    """atan2 of two input dimensions, the sin- and the cos-transformed angle (features.py:213-243): one feature."""

    def __init__(self, idx_sin: int, idx_cos: int):
        if not isinstance(idx_sin, int):
            raise pyrado.TypeErr(given=idx_sin, expected_type=int)
        if not isinstance(idx_cos, int):
            raise pyrado.TypeErr(given=idx_cos, expected_type=int)
        self._idx_sin, self._idx_cos = idx_sin, idx_cos

    def __str__(self):
        return f"ATan2Feat (index for numerator {self._idx_sin}, index for denominator {self._idx_cos})"

    def __call__(self, inp: torch.Tensor) -> torch.Tensor:
        return torch.atan2(inp[..., self._idx_sin], inp[..., self._idx_cos]).unsqueeze(-1)

    def engine_code(self, n_obs: int):
        return ("atan2", [self._idx_sin, self._idx_cos])


class RFFeat:
    """Random Fourier features (Rahimi & Recht 2007; features.py:246-318): `num_feat_per_dim` features
    sqrt(2 / D) cos(freq x + shift) with freq ~ N(0, 1) * sqrt(2 / bandwidth), shift ~ U(0, 2 pi)."""

    def __init__(self, inp_dim: int, num_feat_per_dim: int, bandwidth: Union[float, np.ndarray, torch.Tensor],
                 use_cuda: bool = False):
        self.num_feat_per_dim = num_feat_per_dim
        self.scale = torch.sqrt(torch.tensor(2.0 / num_feat_per_dim))
        self.freq = torch.randn(num_feat_per_dim, inp_dim)
        if not isinstance(bandwidth, torch.Tensor):
            bandwidth = torch.from_numpy(np.asanyarray(bandwidth))
        self.freq *= torch.sqrt(torch.tensor(2.0) / torch.atleast_2d(bandwidth))
        self.shift = 2 * np.pi * torch.rand(num_feat_per_dim)
        self._device = "cuda" if use_cuda and torch.cuda.is_available() else "cpu"
        self.scale, self.freq, self.shift = (t.to(self._device) for t in (self.scale, self.freq, self.shift))

    def __call__(self, inp: torch.Tensor) -> torch.Tensor:
        inp = inp.to(device=self._device, dtype=torch.get_default_dtype())
        if inp.ndimension() > 2:
            raise pyrado.ShapeErr(msg="RBF class can only handle 1-dim or 2-dim input!")
        inp = torch.atleast_2d(inp)
        return self.scale * torch.cos(inp @ self.freq.T + self.shift.repeat(inp.shape[0], 1))

    def engine_code(self, n_obs: int):
        if self.freq.shape[1] != n_obs:
            return None
        return ("rf", [self.num_feat_per_dim, float(self.scale), *self.freq.reshape(-1).tolist(), *self.shift.tolist()])


class RBFFeat:
    """Normalised Gaussian radial basis functions (features.py:321-445): `num_feat_per_dim` centers per input dimension
    on a linspace between the bounds; every dimension's activations sum to one (`state_wise_norm`) or all of them do."""

    def __init__(self, num_feat_per_dim: int, bounds, scale: Union[float, torch.Tensor] = None, state_wise_norm: bool = True,
                 use_cuda: bool = False):
        if not num_feat_per_dim > 1:
            raise pyrado.ValueErr(given=num_feat_per_dim, g_constraint="1")
        if not len(bounds) == 2:
            raise pyrado.ShapeErr(given=bounds, expected_match=np.empty(2))
        b = []
        for x in bounds:
            if isinstance(x, np.ndarray):
                b.append(torch.from_numpy(x))
            elif isinstance(x, torch.Tensor):
                b.append(x.clone())
            elif isinstance(x, (list, tuple)):
                b.append(torch.tensor(x, dtype=torch.get_default_dtype()))
            elif isinstance(x, (int, float)):
                b.append(torch.tensor(x, dtype=torch.get_default_dtype()).view(1))
            else:
                raise pyrado.TypeErr(given=x, expected_type=[np.ndarray, torch.Tensor, list, tuple, int, float])
        if any(bool(torch.isinf(x).any()) for x in b):
            b = [torch.clamp(x, min=-1e6, max=1e6) for x in b]  # the reference clips infinite bounds the same way
        lo, up = b
        num_dim = len(lo)
        self.num_feat = num_feat_per_dim * num_dim
        self.centers = torch.empty(num_feat_per_dim, num_dim)
        for i in range(num_dim):
            self.centers[:, i] = torch.linspace(float(lo[i]), float(up[i]), num_feat_per_dim)
        if scale is None:
            delta = self.centers[1, :] - self.centers[0, :]
            self.scale = -torch.log(torch.tensor(0.2)) / torch.pow(delta, 2)
        else:
            self.scale = torch.as_tensor(scale, dtype=torch.get_default_dtype())
        self._state_wise_norm = state_wise_norm
        self._device = "cuda" if use_cuda and torch.cuda.is_available() else "cpu"
        self.centers, self.scale = self.centers.to(self._device), self.scale.to(self._device)

    @staticmethod
    def _l1(x: torch.Tensor, dim: int) -> torch.Tensor:
        n = torch.norm(x, p=1, dim=dim)
        return x / torch.where(n > 1e-8, n, torch.ones_like(n)).unsqueeze(dim)

    def __call__(self, inp: torch.Tensor) -> torch.Tensor:
        inp = inp.to(device=self._device, dtype=torch.get_default_dtype())
        if inp.ndimension() > 2:
            raise pyrado.ShapeErr(msg="RBF class can only handle 1-dim or 2-dim input!")
        inp = torch.atleast_2d(inp)
        esd = torch.exp(-self.scale * torch.pow(inp.unsqueeze(1) - self.centers, 2))  # [batch, F, dim]
        if self._state_wise_norm:
            return self._l1(esd, 1).transpose(1, 2).reshape(inp.shape[0], -1)
        return self._l1(esd.transpose(1, 2).reshape(inp.shape[0], -1), -1)

    def engine_code(self, n_obs: int):
        if self.centers.shape[1] != n_obs:
            return None
        sc = self.scale.reshape(-1)
        sc = sc.expand(n_obs) if sc.numel() == 1 else sc
        return ("rbf", [self.centers.shape[0], float(self._state_wise_norm), *sc.tolist(), *self.centers.reshape(-1).tolist()])


# feature function -> SRE_FEAT_* name of the C-ABI
ENGINE_FEATS = {const_feat: "const", identity_feat: "identity", sign_feat: "sign", abs_feat: "abs",
                squared_feat: "squared", cubic_feat: "cubic", sin_feat: "sin", cos_feat: "cos",
                sinsin_feat: "sinsin", sincos_feat: "sincos"}


class FeatureStack:
    """Features are nonlinear transformations of the (1-dim) inputs, concatenated along the last dimension."""

    def __init__(self, *feat_fcns: Sequence[Callable[[torch.Tensor], torch.Tensor]]):
        self.feat_fcns = tuple(feat_fcns)
        if len(self.feat_fcns) < 1:
            raise pyrado.ShapeErr(given=self.feat_fcns, expected_match=(-1,))

    def __str__(self):
        return "FeatureStack [" + ", ".join(getattr(f, "__name__", str(f)) for f in self.feat_fcns) + "]"

    def __call__(self, inp: torch.Tensor) -> torch.Tensor:
        vals = [fcn(inp) for fcn in self.feat_fcns]
        if inp.ndim == 1:  # RFFeat / RBFFeat answer a 1-dim input with [1, n]: flattened, so that they stack with the others
            vals = [v.reshape(-1) for v in vals]
        return torch.cat(vals, dim=-1).to(dtype=inp.dtype)

    def get_num_feat(self, inp_flat_dim: int) -> int:
        """Number of feature values: `inp_flat_dim` per element-wise function, one for const / MultFeat / ATan2Feat,
        `num_feat_per_dim` for RFFeat, `num_feat` for RBFFeat (the reference's bookkeeping, features.py:77-128, gives the
        same for the stacks it supports; it miscounts stacks that mix RFFeat / RBFFeat with other functions)."""
        n = 0
        for f in self.feat_fcns:
            if f is const_feat or isinstance(f, (MultFeat, ATan2Feat)):
                n += 1
            elif isinstance(f, RFFeat):
                n += f.num_feat_per_dim
            elif isinstance(f, RBFFeat):
                n += f.num_feat
            else:
                n += inp_flat_dim
        return n

    def engine_codes(self, n_obs: Optional[int] = None) -> Optional[list]:
        """The stack as SRE_FEAT_* names -- (name, parameters) for the parametrised features --, or None if it holds a
        function the kernel does not implement.  `functools.partial(sig_feat, scale=..)` / `bell_feat` carry their scale."""
        out = []
        for f in self.feat_fcns:
            if f in ENGINE_FEATS:
                out.append(ENGINE_FEATS[f])
            elif f is sig_feat or f is bell_feat:
                out.append(("sig" if f is sig_feat else "bell", [1.0]))
            elif isinstance(f, functools.partial) and f.func in (sig_feat, bell_feat) and not f.args and set(f.keywords) <= {"scale"}:
                out.append(("sig" if f.func is sig_feat else "bell", [float(f.keywords.get("scale", 1.0))]))
            elif hasattr(f, "engine_code") and n_obs is not None and f.engine_code(n_obs) is not None:
                out.append(f.engine_code(n_obs))
            else:
                return None
        return out
