"""lya_cands.py:72-94 of the reference: the DecisionBoundary Lyapunov candidate.  Stand-alone it is one small kernel
(fiode_decision_boundary); inside the FI loss it is fused into the loss kernel (fiode_fi_loss_fwd), which covers
on_simplex=True, log_mode=False (the README configs)."""
from torch import nn

from . import ops


class DecisionBoundary(nn.Module):
    def __init__(self, on_simplex=False, log_mode=False, num_class=10):
        super().__init__()
        assert num_class == 10, 'the FI-ODE CIFAR configs use 10 classes'
        self.on_simplex, self.log_mode, self.num_class = on_simplex, log_mode, num_class

    def forward(self, state_output, y):
        """V = 1 + max_{j != y} p_j - p_y (log of it in log_mode), p = state_output or its softmax."""
        return ops.decision_boundary(state_output, y, self.on_simplex, self.log_mode)
