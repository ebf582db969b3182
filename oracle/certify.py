"""oracle/certify.py -- TEST INFRASTRUCTURE ONLY.

CPU restatement of the Lipschitz grid certification and the decision-boundary grid.

Follows:
  * ``robustness/certify_lipschitz.py:37-42``   h_vdot
  * ``robustness/certify_lipschitz.py:65-83``   constants (Lfx, kappa, dist)
  * ``robustness/certify_lipschitz.py:100-143`` chunking, Lf_eta, perturb, tie mask, verdict
  * ``robustness/eval_utils.py:31-89``          sample_decision_boundary / get_grid_for_label /
                                                count_samples_decision_boundary
"""
from __future__ import annotations

import math
import numpy as np
import torch

from .dynamics import eval_dot, dyn_Q


# ---- grid ---------------------------------------------------------------------
def count_boundary_points(n=10, T=40):
    """eval_utils.py:72-89 restated as the closed combinatorial count of
    {z in Z^n_{>=0}: sum z = T, z_0 = max_{i>=1} z_i}: sum over the shared maximum m and the
    number r>=1 of runner-up slots that attain it."""
    total = 0
    for m in range(0, T + 1):
        rest = T - m                       # mass on the n-1 wrong slots, each <= m, max == m
        total += _bounded_compositions(n - 1, rest, m) - (_bounded_compositions(n - 1, rest, m - 1) if m > 0 else 0)
    return total


def _bounded_compositions(k, total, cap):
    """#{x in Z^k, 0<=x_i<=cap, sum x = total} by inclusion-exclusion."""
    if cap < 0:
        return 0
    out = 0
    for j in range(0, k + 1):
        rem = total - j * (cap + 1)
        if rem < 0:
            break
        out += (-1) ** j * math.comb(k, j) * math.comb(rem + k - 1, k - 1)
    return out


def boundary_grid_numerators(n=10, T=10):
    """Integer numerators z (N, n) of the label-0 grid: z_0 = max_{i>=1} z_i, sum z = T
    (the set eval_utils.py:31-61 builds; order is canonical-lexicographic here, the
    reference's DP order differs -- every consumer is a max-reduction over rows)."""
    rows = []

    def rec(prefix, remaining, slots, cap, hit):
        if slots == 0:
            if remaining == 0 and hit:
                rows.append(prefix)
            return
        lo = max(0, remaining - cap * (slots - 1))
        for zi in range(lo, min(cap, remaining) + 1):
            rec(prefix + [zi], remaining - zi, slots - 1, cap, hit or zi == cap)

    for m in range(0, T + 1):
        if T - m < m or (n - 1) * m < T - m:
            continue
        rec([m], T - m, n - 1, m, False)
    if not rows:
        return np.zeros((0, n), dtype=np.int64)
    return np.asarray(rows, dtype=np.int64)


def boundary_grid_numerators_c(n=10, T=40):
    """Same rows in the same order as boundary_grid_numerators, from the plain-C restatement
    oracle/grid_enum.c (uint8; 41 320 837 x 10 at T = 40 in about a second)."""
    import ctypes
    from . import build as _b
    lib = ctypes.CDLL(str(_b.build()))
    lib.oracle_boundary_grid.restype = ctypes.c_longlong
    lib.oracle_boundary_grid.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    rows = lib.oracle_boundary_grid(n, T, None)
    assert rows >= 0, 'need 2 <= n <= 16, 1 <= T <= 255'
    out = np.empty((rows, n), dtype=np.uint8)
    got = lib.oracle_boundary_grid(n, T, out.ctypes.data_as(ctypes.c_void_p))
    assert got == rows
    return out


def grid_row_hash(z):
    """Order-sensitive 64-bit checksum of a (N, n) numerator array: sum_i (i + 1) * key_i mod 2^64 with key_i the
    base-256 packing of row i -- lets a test compare two 41 M-row enumerations without holding both."""
    z = np.ascontiguousarray(z, dtype=np.uint8)
    key = np.zeros(z.shape[0], dtype=np.uint64)
    for j in range(min(z.shape[1], 8)):
        key = key * np.uint64(256) + z[:, j].astype(np.uint64)
    for j in range(8, z.shape[1]):
        key = key ^ (z[:, j].astype(np.uint64) << np.uint64(8 * (j - 8) + 5))
    idx = np.arange(1, z.shape[0] + 1, dtype=np.uint64)
    with np.errstate(over='ignore'):
        return int((key * idx).sum(dtype=np.uint64))


def grid_for_label(z, T, label):
    """eval_utils.py:64-69: swap columns 0 and ``label``; float32 grid = z / T."""
    g = z.astype(np.float64) / T
    if label != 0:
        g[:, [label, 0]] = g[:, [0, label]]
    return torch.from_numpy(g).float()


# ---- certification ------------------------------------------------------------
def certify_constants(cfg, std_min, eps, T, n_class=10):
    """certify_lipschitz.py:65-83."""
    Lfx = (cfg['alpha_1'] if cfg['scale_nominal'] else 1.0) / std_min
    kappa = math.sqrt(2) * Lfx * eps
    dist = math.sqrt(n_class) / T
    return dict(Lfx=Lfx, kappa=kappa, dist=dist, eps_grid=1.0 / T)


def h_vdot(eta, s, label, runner_up, sd, cfg, Q, prefix):
    """certify_lipschitz.py:37-42 with eval_dot_light (classification.py:117-126)."""
    f = eval_dot(eta, s, sd, cfg, Q, prefix)
    f_y = f[:, label]
    f_wrong = f.masked_fill(~runner_up, -float('inf')).max(dim=-1).values
    return -f_y + f_wrong


@torch.no_grad()
def certify_image(grid_label, s, label, sd, cfg, eps=0.141, T=40, batches=10, std_min=0.225,
                  prefix='model.dyn_fun.', Q=None, n_class=10):
    """certify_lipschitz.py:100-143 for ONE image.  grid_label: (N_g, 10) float32 grid already
    column-swapped for ``label``; s: (1, 10) static feature.  Returns
    (max_violation, max_violation_larger_T) -- the image is certified iff max_violation < 0."""
    c = certify_constants(cfg, std_min, eps, T, n_class)
    Q = Q or dyn_Q(sd, prefix)
    n = grid_label.shape[0]
    chunk = n // batches
    nb = batches + (1 if n % batches != 0 else 0)
    viol, viol_T = [], []
    for bi in range(nb):
        if (bi + 1) * chunk < n:
            eta = grid_label[bi * chunk:(bi + 1) * chunk]
        else:
            eta = grid_label[bi * chunk:]
        if eta.shape[0] == 0:
            continue
        eta_ub = eta.amax(dim=1) + c['eps_grid']
        Lf_eta = math.sqrt(n_class) * (cfg['sigma_1'] * cfg['alpha_1'] * torch.exp(cfg['sigma_1'] * eta_ub)) + 1
        perturb = (math.sqrt(2) * Lf_eta) * c['dist']
        ind_wrong = eta == eta.max(dim=-1, keepdim=True).values
        ind_wrong[:, label] = False
        hx = h_vdot(eta, s, label, ind_wrong, sd, cfg, Q, prefix)
        viol_T.append((hx + c['kappa']).max().item())
        viol.append((hx + perturb + c['kappa']).max().item())
    return max(viol), max(viol_T)
