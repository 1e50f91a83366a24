"""
FNN and FNNPolicy (pyrado/policies/feed_back/fnn.py:43-219).  The torch modules serve the learner; sampling uses the
fused CUDA forward pass: tanh/relu/identity hidden layers, optional tanh output (csrc/sre_policy.cuh).  The hot shapes
[obs, H, H, act] with H in {16, 32, 64} and tanh run on the warp-tiled packed-FMA kernel.
"""
from typing import Callable, Iterable, Optional, Sequence, Union

import torch
import torch.nn as nn
from torch.nn.utils import convert_parameters as cp

from ... import _lib as L
from ... import engine as E
from ...utils.data_types import EnvSpec
from ..base import Policy
from ..initialization import init_param

_NONLIN_CODES = {torch.tanh: 0, torch.relu: 1, nn.functional.relu: 1, None: 2}


def _nonlin_code(fcn) -> Optional[int]:
    if isinstance(fcn, nn.Tanh):
        return 0
    if isinstance(fcn, nn.ReLU):
        return 1
    try:
        return _NONLIN_CODES.get(fcn)
    except TypeError:
        return None


class FNN(nn.Module):
    """Feed-forward neural network"""

    def __init__(self, input_size: int, output_size: int, hidden_sizes: Sequence[int],
                 hidden_nonlin: Union[Callable, Sequence[Callable]], dropout: Optional[float] = 0.0,
                 output_nonlin: Optional[Callable] = None, init_param_kwargs: Optional[dict] = None,
                 use_cuda: bool = False):
        self._device = "cuda" if use_cuda and torch.cuda.is_available() else "cpu"
        super().__init__()
        self.hidden_nonlin = (hidden_nonlin if isinstance(hidden_nonlin, Iterable) else len(hidden_sizes) * [hidden_nonlin])
        self.dropout = dropout
        self.output_nonlin = output_nonlin
        self.sizes = [input_size, *hidden_sizes, output_size]
        self.hidden_layers = nn.ModuleList()
        last = input_size
        for hs in hidden_sizes:
            self.hidden_layers.append(nn.Linear(last, hs))
            last = hs
            if self.dropout > 0:
                self.hidden_layers.append(nn.Dropout(p=self.dropout))
        self.output_layer = nn.Linear(last, output_size)
        self.init_param(None, **(init_param_kwargs or dict()))
        self.to(self.device)

    @property
    def device(self) -> str:
        return self._device

    @property
    def param_values(self) -> torch.Tensor:
        return cp.parameters_to_vector(self.parameters())

    @param_values.setter
    def param_values(self, param: torch.Tensor):
        cp.vector_to_parameters(param, self.parameters())

    def init_param(self, init_values: Optional[torch.Tensor] = None, **kwargs):
        if init_values is None:
            for layer in self.hidden_layers:
                if isinstance(layer, nn.Linear):
                    init_param(layer, **kwargs)
            init_param(self.output_layer, **kwargs)
        else:
            self.param_values = init_values

    def forward(self, obs: torch.Tensor) -> torch.Tensor:
        x = obs.to(device=self.device)
        i_lin = 0
        for layer in self.hidden_layers:
            x = layer(x)
            if isinstance(layer, nn.Linear):
                if self.hidden_nonlin[i_lin] is not None:
                    x = self.hidden_nonlin[i_lin](x)
                i_lin += 1
        x = self.output_layer(x)
        return self.output_nonlin(x) if self.output_nonlin is not None else x

    def engine_codes(self):
        """(hidden_nonlin, output_nonlin) codes of the C-ABI, or None if the kernels do not implement this net."""
        codes = {_nonlin_code(f) for f in self.hidden_nonlin}
        out = 0 if self.output_nonlin is None else (1 if _nonlin_code(self.output_nonlin) == 0 else None)
        if len(codes) != 1 or None in codes or out is None or self.dropout > 0:
            return None
        if len(self.sizes) - 1 > L.MAX_LAYERS or max(self.sizes) > L.MAX_WIDTH:
            return None
        return codes.pop(), out


class FNNPolicy(Policy):
    """Feed-forward neural network policy"""

    name: str = "fnn"

    def __init__(self, spec: EnvSpec, hidden_sizes: Sequence[int], hidden_nonlin: Union[Callable, Sequence[Callable]],
                 dropout: Optional[float] = 0.0, output_nonlin: Optional[Callable] = None,
                 init_param_kwargs: Optional[dict] = None, use_cuda: bool = False, cartpole_feats: bool = False):
        """
        :param cartpole_feats: the fork's input transform [x, th, xd, thd] -> [x, sin th, cos th, xd, thd] applied
                               inside `forward` with `obs_dim + 1` network inputs (fnn.py:193, :214-218); the default
                               is the upstream behaviour, `net(obs)` with `obs_dim` inputs
        """
        super().__init__(spec, use_cuda)
        self.cartpole_feats = bool(cartpole_feats)
        self.net = FNN(input_size=spec.obs_space.flat_dim + int(self.cartpole_feats), output_size=spec.act_space.flat_dim,
                       hidden_sizes=hidden_sizes, hidden_nonlin=hidden_nonlin, dropout=dropout,
                       output_nonlin=output_nonlin, use_cuda=use_cuda)
        self.init_param(None, **(init_param_kwargs or dict()))
        self.to(self.device)

    def init_param(self, init_values: torch.Tensor = None, **kwargs):
        if init_values is None:
            self.net.init_param(init_values, **kwargs)
        else:
            self.param_values = init_values

    def forward(self, obs: torch.Tensor) -> torch.Tensor:
        if self.cartpole_feats:
            obs = torch.cat([obs[..., 0:1], torch.sin(obs[..., 1:2]), torch.cos(obs[..., 1:2]), obs[..., 2:]], dim=-1)
        return self.net(obs)

    def engine_desc(self, device, param_sets=None, group_size: int = 0):
        codes = self.net.engine_codes()
        if codes is None:
            return None
        weights = self._engine_weights(device) if param_sets is None else param_sets
        host = getattr(self, "_host_weights", None) if param_sets is None else None  # small layers travel as launch arguments
        return E.make_policy(L.POL_FNN, weights, sizes=self.net.sizes, hidden_nonlin=codes[0],
                             output_nonlin=codes[1], cartpole_feats=self.cartpole_feats,
                             precision=self.engine_precision, group_size=group_size, weights_host=host)
