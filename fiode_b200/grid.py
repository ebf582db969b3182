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


# ---- grid file format ---------------------------------------------------------------
def save_grid_pt(path, n=10, T=40, device='cuda'):
    """Write the file ``robustness/sample_decision_boundary.py:37-42`` writes (``grid_T.pt``): a Python list of n
    float32 CPU tensors (N, n), entry ``i`` being the label-0 grid with columns 0 and i swapped
    (eval_utils.py:64-69, ``.float()`` of the float64 z / T).  The reference builds it with an O(N^2) numpy
    recursion (minutes and > 10 GB at T = 40); here the numerators are enumerated on the device and each label's
    tensor is produced by one column swap.  ``certify_lipschitz.py:57-58`` loads it with ``torch.load``."""
    z = boundary_grid_numerators(n, T, device)
    base = (z.double() / T).float()                     # same rounding as th.from_numpy(float64).float()
    grids = [get_grid_for_label(base, i).cpu() for i in range(n)]
    torch.save(grids, path)
    return grids


def numerators_from_grid_pt(path_or_list, T, label=0, device='cuda'):
    """Read a reference ``grid_T.pt`` (or the list it holds) back into the uint8 numerators of the canonical
    label-0 grid that ``fiode_certify_grid`` streams: z = round(T * eta), columns un-swapped.  Checks that every row
    sums to T and lies on the decision boundary (z_0 = max_{i>=1} z_i, eval_utils.py:31-61)."""
    grids = torch.load(path_or_list, map_location='cpu') if isinstance(path_or_list, (str, bytes)) or hasattr(path_or_list, '__fspath__') else path_or_list
    g = grids[label].to(torch.float64)
    if label != 0:
        g = g.clone()
        g[:, [label, 0]] = g[:, [0, label]]
    z = torch.round(g * T).to(torch.int64)
    if not bool((z.sum(1) == T).all()):
        raise ValueError('grid rows do not sum to T')
    if not bool((z[:, 0] == z[:, 1:].max(1).values).all()):
        raise ValueError('grid rows are not on the label-0 decision boundary')
    if int(z.max()) > 255:
        raise ValueError('T too large for uint8 numerators')
    return z.to(torch.uint8).to(device)
