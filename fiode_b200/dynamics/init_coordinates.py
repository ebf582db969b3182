"""dynamics/init_coordinates.py:4-44 of the reference (host-side bookkeeping only)."""
import torch
from torch import nn


class DefaultInitFun(nn.Module):
    def __init__(self, h_dims, param_map=nn.Identity()):
        super().__init__()
        self.h_dims = tuple(h_dims)
        self.param_map = param_map
        for i, d in enumerate(self.h_dims):
            self.register_buffer(f'h0_{i}', self._h0(d))

    @staticmethod
    def _h0(d):
        return torch.zeros((d,))

    @property
    def h0s(self):
        return [getattr(self, f'h0_{i}') for i in range(len(self.h_dims))]

    def update_prior(self, h_prior):
        assert len(h_prior) == len(self.h_dims), '[ERROR] Inconsistent attribute lengths'
        for i in range(len(self.h_dims)):
            setattr(self, f'h0_{i}', h_prior[i])

    def forward(self, x, dyn):
        if callable(getattr(dyn, 'state_init', None)):
            state = tuple(dyn.state_init(x))
        elif self.h0s[0].ndim == 2:
            state = tuple(h.clone().to(x.device) for h in self.h0s)
        else:
            state = tuple(h.clone()[None, :].repeat(x.shape[0], 1).to(x.device) for h in self.h0s)
        return self.param_map(x), state


class UniformInitFun(DefaultInitFun):
    @staticmethod
    def _h0(d):
        return torch.ones((d,)) / d
