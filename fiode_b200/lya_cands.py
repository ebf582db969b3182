"""lya_cands.py:72-94 of the reference: the DecisionBoundary Lyapunov candidate.  Stand-alone it is a
tiny reduction; inside the FI loss it is fused into the loss kernel (fiode_fi_loss_fwd)."""
import torch
from torch import nn

from . import ops


class DecisionBoundary(nn.Module):
    def __init__(self, on_simplex=False, log_mode=False, num_class=10):
        super().__init__()
        assert on_simplex and num_class == 10, 'the fused FI loss covers on_simplex=True, 10 classes (README configs)'
        self.on_simplex, self.log_mode, self.num_class = on_simplex, log_mode, num_class

    def forward(self, state_output, y):
        """V = 1 + max_{j != y} p_j - p_y, computed by the FI-loss kernel with a zero static feature
        (V does not depend on f)."""
        N = state_output.shape[0]
        dyn = getattr(self, '_dyn_for_v', None)
        assert dyn is not None, 'attach a dynamics module (LyapunovLearning does) before calling DecisionBoundary alone'
        s = torch.zeros(N, 10, device=state_output.device)
        sv = ops.fi_loss_fwd(dyn.dyn_weights(), state_output, s, y, 1, 0.0)
        return torch.log(sv.V) if self.log_mode else sv.V
