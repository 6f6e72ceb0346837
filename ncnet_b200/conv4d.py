"""Drop-in for the reference's ``lib/conv4d.py``: ``conv4d`` and ``Conv4d`` with the same
signatures, state_dict layout and attributes, running on this library's CUDA kernels."""
from __future__ import annotations

import math

import torch
from torch.nn import Module
from torch.nn.parameter import Parameter

from .functional import conv4d  # noqa: F401  (re-exported: same name and arguments as lib/conv4d.py:11)


class Conv4d(Module):
    """4D convolution layer (reference lib/conv4d.py:53-82).

    ``weight`` is stored the way the reference stores it: created as [Cout,Cin,k,k,k,k] with the
    default ``_ConvNd`` initialisation and, for ``pre_permuted_filters=True`` (default), kept as
    [k,Cout,Cin,k,k,k] -- the shape found in the published checkpoints.  ``use_half`` is kept as
    an attribute because ``ImMatchNet`` toggles it (lib/model.py:256-258); the kernels take the
    storage type from the input tensor and always accumulate in fp32.
    """

    def __init__(self, in_channels, out_channels, kernel_size, bias=True, pre_permuted_filters=True):
        super().__init__()
        if isinstance(kernel_size, (tuple, list)):
            if len(set(kernel_size)) != 1:
                raise ValueError("Conv4d: only cubic kernels are supported")
            kernel_size = kernel_size[0]
        k = int(kernel_size)
        if k not in (3, 5):
            # the reference accepts any odd size; this library's kernels cover the sizes its configurations use
            raise ValueError(f"Conv4d: kernel_size {k} is outside this library's envelope (3 or 5)")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size = (k, k, k, k)
        self.stride, self.padding, self.dilation, self.groups = (1,) * 4, (0,) * 4, (1,) * 4, 1
        self.pre_permuted_filters = pre_permuted_filters
        w = torch.empty(out_channels, in_channels, k, k, k, k)
        torch.nn.init.kaiming_uniform_(w, a=math.sqrt(5))          # what _ConvNd.reset_parameters does
        bound = 1.0 / math.sqrt(in_channels * k ** 4)
        if pre_permuted_filters:
            w = w.permute(2, 0, 1, 3, 4, 5).contiguous()
        self.weight = Parameter(w)
        if bias:
            self.bias = Parameter(torch.empty(out_channels).uniform_(-bound, bound))
        else:
            self.register_parameter("bias", None)
        self.use_half = False

    def forward(self, input):
        if self.out_channels > 16 and torch.is_grad_enabled() and self.weight.requires_grad:
            raise RuntimeError(f"Conv4d: the weight gradient covers at most 16 output channels (this layer has "
                               f"{self.out_channels}); run it under torch.no_grad() or freeze its weight")
        return conv4d(input, self.weight, bias=self.bias, permute_filters=not self.pre_permuted_filters,
                      use_half=self.use_half)

    def extra_repr(self):
        return f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}"
