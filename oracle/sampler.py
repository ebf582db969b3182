"""oracle/sampler.py -- TEST INFRASTRUCTURE ONLY.

CPU restatement of the simplex samplers, as pure functions of pre-drawn Exp(1) variates
(the reference draws them with ``torch.distributions.Exponential.sample``; parity tests feed
the same variates to both sides).

Follows ``sampling/sampler.py``: UniformSimplexSampling ``:24-38``, CorrectConeSampling
``:104-128``, DecisionBoundarySampling ``:130-153``, CompositeSampler ``:169-216``.
"""
from __future__ import annotations

from math import floor
import torch
import torch.nn.functional as F


def uniform_simplex(e):
    """e: (n, h) Exp(1) -> L1-normalised rows (one set shared by every image)."""
    return F.normalize(e, p=1.0, dim=1)


def correct_cone(e, y):
    """e: (B, n, h) Exp(1), y: (B,) -> uniform simplex samples with the row maximum swapped
    into the label slot (sampler.py:113-125)."""
    h = F.normalize(e, dim=-1, p=1)
    hmax, imax = h.max(dim=-1)
    yy = y[:, None].expand(-1, h.shape[1])
    hlabel = h.gather(2, yy[:, :, None])[:, :, 0]
    out = h.clone()
    out.scatter_(2, yy[:, :, None], hmax[:, :, None])
    out.scatter_(2, imax[:, :, None], hlabel[:, :, None])
    return out


def decision_boundary(e, y):
    """e: (B, n, h-1) Exp(1): z1 = max(zs); [z1, zs]/sum; label slot <- z1, the other slots
    keep zs in order (sampler.py:139-150)."""
    B, n, hm1 = e.shape
    z1 = e.max(dim=-1).values[:, :, None]
    raw = F.normalize(torch.cat([z1, e], dim=-1), p=1, dim=-1)
    out = torch.zeros(B, n, hm1 + 1, dtype=e.dtype)
    for b in range(B):
        cols = [j for j in range(hm1 + 1) if j != int(y[b])]
        out[b, :, int(y[b])] = raw[b, :, 0]
        out[b][:, cols] = raw[b, :, 1:]
    return out


def split_counts(sample_size, coeffs):
    """sampler.py:181-192."""
    out, added = [], 0
    for i, c in enumerate(coeffs):
        if i == len(coeffs) - 1:
            out.append(sample_size - added)
            break
        k = floor(sample_size * c)
        added += k
        out.append(k)
    assert sum(out) == sample_size
    return out


def composite(parts, batch_size):
    """sampler.py:203-216: 2-D parts are repeated per image, then cat on the sample axis and
    flattened to (B*S, h)."""
    full = [p[None].repeat(batch_size, 1, 1) if p.ndim == 2 else p for p in parts]
    return torch.cat(full, dim=1).flatten(0, 1)
