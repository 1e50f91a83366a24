"""
FeatureStack and the element-wise feature functions (pyrado/policies/features.py:42-175).  The functions below are
the ones the Linear-policy kernel implements (csrc/sre_policy.cuh ``feat_f``); a stack holding any other callable
still works as a torch module but is evaluated outside the fused kernel.
"""
from typing import Callable, Optional, Sequence

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
        return torch.cat([fcn(inp) for fcn in self.feat_fcns], dim=-1).to(dtype=inp.dtype)

    def get_num_feat(self, inp_flat_dim: int) -> int:
        """Every function yields `inp_flat_dim` features, the constant feature one."""
        return sum(1 if f is const_feat else inp_flat_dim for f in self.feat_fcns)

    def engine_codes(self) -> Optional[list]:
        """The stack as SRE_FEAT_* names, or None if it holds a function the kernel does not implement."""
        if all(f in ENGINE_FEATS for f in self.feat_fcns):
            return [ENGINE_FEATS[f] for f in self.feat_fcns]
        return None
