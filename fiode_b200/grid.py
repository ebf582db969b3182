"""Decision-boundary grid (robustness/eval_utils.py:31-89 of the reference) enumerated on the GPU."""
from __future__ import annotations

import torch

from ._lib import lib, check, ptr, stream_ptr


def count_samples_decision_boundary(n=10, T=40) -> int:
    """eval_utils.py:72-89 (host-side table, no GPU needed)."""
    c = int(lib().fiode_grid_count(n, T))
    if c < 0:
        check(-1, 'fiode_grid_count')
    return c


def boundary_grid_numerators(n=10, T=40, device='cuda', first=0, count=None):
    """uint8 numerators z (N, n) of the canonical label-0 grid eta = z / T."""
    total = count_samples_decision_boundary(n, T)
    count = total - first if count is None else count
    out = torch.empty(count, n, dtype=torch.uint8, device=device)
    nb = lib().fiode_grid_workspace_bytes(n, T)
    ws = torch.empty(nb, dtype=torch.uint8, device=device)
    with torch.cuda.device(out.device):
        check(lib().fiode_grid_enumerate(n, T, first, count, ptr(out), ptr(ws), nb, stream_ptr()), 'fiode_grid_enumerate')
    return out


def sample_decision_boundary(n=10, T=40, device='cuda'):
    """eval_utils.py:31-61: float64 grid eta = z / T for label 0 (row order is canonical, not the
    reference's DP order; every consumer is a max-reduction over rows)."""
    return boundary_grid_numerators(n, T, device).double() / T


def get_grid_for_label(grid, label):
    """eval_utils.py:64-69: swap columns 0 and ``label``."""
    g = grid.clone()
    if label != 0:
        g[:, [label, 0]] = g[:, [0, label]]
    return g
