"""oracle/layers.py -- TEST INFRASTRUCTURE ONLY.

CPU (torch fp32 / complex64) restatement of the Cayley-orthogonal layers and the
KWLarge_Concat backbone.  Works on a plain ``dict`` of tensors keyed like the
reference state_dict -- no nn.Module involved.

Follows:
  * ``libs/ortho_conv/layers.py:50-60``   cayley
  * ``libs/ortho_conv/layers.py:24-48``   StridedConv (stride-2 emulation, space-to-depth)
  * ``libs/ortho_conv/layers.py:62-86``   CayleyConv.forward
  * ``libs/ortho_conv/layers.py:88-105``  CayleyLinear.forward
  * ``libs/ortho_conv/layers.py:127-131`` GroupSort
  * ``libs/ortho_conv/models_ortho.py:3-28`` KWLarge_Concat
  * ``models.py:17-35`` Normalize, make_ortho_KWLarge_Concat
"""
from __future__ import annotations

import math
import torch
import torch.nn.functional as F


def cayley(W: torch.Tensor) -> torch.Tensor:
    """layers.py:50-60.  W (..., cout, cin) real or complex -> (semi-)orthogonal Q."""
    if W.dim() == 2:
        return cayley(W.unsqueeze(0)).squeeze(0)
    cout, cin = W.shape[-2:]
    if cin > cout:                       # plain (non-conjugate) transpose, layers.py:54-55
        return cayley(W.transpose(1, 2)).transpose(1, 2)
    U, V = W[:, :cin], W[:, cin:]
    eye = torch.eye(cin, dtype=W.dtype).unsqueeze(0)
    A = U - U.conj().transpose(1, 2) + V.conj().transpose(1, 2) @ V
    M = torch.inverse(eye + A)
    return torch.cat((M @ (eye - A), -2 * V @ M), dim=1)


def space_to_depth(x: torch.Tensor) -> torch.Tensor:
    """layers.py:45-48: 'b c (w k1) (h k2) -> b (c k1 k2) w h', k1=k2=2."""
    b, c, W2, H2 = x.shape
    w, h = W2 // 2, H2 // 2
    return x.view(b, c, w, 2, h, 2).permute(0, 1, 3, 5, 2, 4).reshape(b, c * 4, w, h)


def conv_weight_spectrum(weight: torch.Tensor, n: int) -> torch.Tensor:
    """layers.py:74-79: shifted, conjugated half-spectrum of the kernel, (F, cout, cin)."""
    cout, cin, k, _ = weight.shape
    s = (k - 1) // 2
    idx = torch.arange(0, n).repeat((n, 1))
    idx = idx + idx.T
    shift = torch.exp(1j * 2 * math.pi * (-s) * idx / n)[:, : n // 2 + 1].reshape(n * (n // 2 + 1), 1, 1)
    wf = torch.fft.rfft2(weight, (n, n)).reshape(cout, cin, n * (n // 2 + 1)).permute(2, 0, 1).conj()
    return shift * wf


def conv_alpha_init(weight: torch.Tensor, n: int) -> float:
    """layers.py:80-81: the lazily created alpha = ||wfft||_F over the half spectrum."""
    return conv_weight_spectrum(weight, n).norm().item()


def cayley_conv_Q(weight, alpha, n):
    wf = conv_weight_spectrum(weight, n)
    return cayley(alpha * wf / wf.norm())


def cayley_conv(x, weight, bias, alpha, stride=1):
    """layers.py:72-86 (+ the stride-2 pre-hook).  ``weight`` is the stored kernel:
    (cout, cin, 3, 3) for stride 1, (cout, 4*cin, 1, 1) for the emulated stride 2."""
    if stride == 2:
        x = space_to_depth(x)
    cout, cin = weight.shape[:2]
    b, _, n, _ = x.shape
    Fq = n * (n // 2 + 1)
    xf = torch.fft.rfft2(x).permute(2, 3, 1, 0).reshape(Fq, cin, b)
    yf = (cayley_conv_Q(weight, alpha, n) @ xf).reshape(n, n // 2 + 1, cout, b)
    y = torch.fft.irfft2(yf.permute(3, 2, 0, 1))
    if bias is not None:
        y = y + bias[:, None, None]
    return y


def cayley_linear_Q(weight, alpha):
    """layers.py:103-104."""
    return cayley(alpha * weight / weight.norm())


def cayley_linear(x, weight, bias, alpha):
    return F.linear(x, cayley_linear_Q(weight, alpha), bias)


def group_sort(x):
    """layers.py:127-131: channel halves -> (max, min)."""
    a, b = x.split(x.size(1) // 2, 1)
    return torch.cat([torch.max(a, b), torch.min(a, b)], dim=1)


# ---- backbone -----------------------------------------------------------------
CONV_LAYERS = [  # (index in model Sequential, stride, spatial n seen by the layer)
    (0, 1, 32), (2, 2, 16), (4, 1, 16), (6, 2, 8)]
LIN_LAYERS = [9, 11]


def backbone_forward(x, sd, prefix=''):
    """Normalize -> KWLarge_Concat(out_dim=10, act=GroupSort).  ``sd`` keys, relative to
    ``prefix`` (= 'model.init_coordinates.param_map.' in a Lightning checkpoint):
    '0.mu', '0.std', '1.model.{0,2,4,6,9,11}.{weight,bias,alpha}', '1.fc.{weight,bias,alpha}'.
    models.py:23-35, models_ortho.py:12-28."""
    g = lambda k: sd[prefix + k]
    h = (x - g('0.mu')) / g('0.std')
    for idx, stride, _n in CONV_LAYERS:
        p = f'1.model.{idx}.'
        h = group_sort(cayley_conv(h, g(p + 'weight'), g(p + 'bias'), g(p + 'alpha'), stride))
    h = h.flatten(1)
    for idx in LIN_LAYERS:
        p = f'1.model.{idx}.'
        h = group_sort(cayley_linear(h, g(p + 'weight'), g(p + 'bias'), g(p + 'alpha')))
    return cayley_linear(h, g('1.fc.weight'), g('1.fc.bias'), g('1.fc.alpha'))
