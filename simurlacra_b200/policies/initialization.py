"""init_param for the layer types on the path (pyrado/policies/initialization.py:48-76): torch's default Linear
initialisation (Kaiming-uniform weights with a = sqrt(5), uniform bias in +-1/sqrt(fan_in)); orthogonal / normal for
bare parameters."""
from math import sqrt

import torch.nn as nn
import torch.nn.init as init


def init_param(m, **kwargs):
    if isinstance(m, nn.Parameter):
        if m.ndim >= 2:
            init.orthogonal_(m.data)
        else:
            init.normal_(m.data)
    elif isinstance(m, nn.Linear):
        init.kaiming_uniform_(m.weight, a=sqrt(5))
        if m.bias is not None:
            fan_in, _ = init._calculate_fan_in_and_fan_out(m.weight)
            bound = 1 / sqrt(fan_in) if fan_in > 0 else 0
            init.uniform_(m.bias, -bound, bound)
    # other layer types (recurrent, convolutional, ...) are not on the pysim path
