"""sampling/sampler.py of the reference (UniformSimplexSampling :24-38, CorrectConeSampling :104-128,
DecisionBoundarySampling :130-153, CompositeSampler :169-216) on libfiode_b200.so.

The Exp(1) variates are drawn with torch's generator on the device -- the same call the reference
makes (torch.distributions.Exponential.sample == empty().exponential_() / rate); the transformation
of the variates into simplex samples and their composition into the (B*S, 10) tensor is one CUDA
kernel per sampler writing straight into the composite buffer.  ``variates=`` lets a caller supply
the draws (parity tests feed the reference's own)."""
from __future__ import annotations

from math import floor

import torch
from torch import nn

from .. import ops


class AbstractSampler(nn.Module):
    kind = None
    variate_dims = 10

    def __init__(self, h_dims):
        super().__init__()
        self.h_dims = h_dims
        self.device = None

    def device_initialize(self, device):
        self.device = device

    def variate_shape(self, n_batch, sample_size, h_dim):
        raise NotImplementedError

    def draw(self, n_batch, sample_size, h_dim, device):
        return torch.empty(self.variate_shape(n_batch, sample_size, h_dim), device=device).exponential_()

    def forward(self, x, y, model, h_dim, sample_size, variates=None):
        """Stand-alone use, same return shape as the reference: (n, h) for the shared uniform set,
        (B, n, h) otherwise."""
        B = x.shape[0]
        e = self.draw(B, sample_size, h_dim, x.device) if variates is None else variates.to(x.device)
        shared = self.kind == 'uniform'
        h = torch.empty(1 if shared else B, sample_size, h_dim, dtype=torch.float32, device=x.device)
        ops.sampler_fill(self.kind, e, None if shared else y, h, 0)
        return h[0] if shared else h


class UniformSimplexSampling(AbstractSampler):
    kind = 'uniform'

    def variate_shape(self, n_batch, sample_size, h_dim):
        return (sample_size, h_dim)


class CorrectConeSampling(AbstractSampler):
    kind = 'correct_cone'

    def variate_shape(self, n_batch, sample_size, h_dim):
        return (n_batch, sample_size, h_dim)


class DecisionBoundarySampling(AbstractSampler):
    kind = 'decision_boundary'

    def variate_shape(self, n_batch, sample_size, h_dim):
        return (n_batch, sample_size, h_dim - 1)


class CompositeSampler(nn.Module):
    def __init__(self, h_dims, samplers):
        super().__init__()
        self.samplers = samplers
        self.h_dims = h_dims

    def device_initialize(self, device):
        for s in self.samplers:
            s.device_initialize(device)

    def _coefficient_to_num_samples(self, sample_size, mixer_coefficients):
        mixed, added = [], 0
        for coeff in mixer_coefficients:
            if len(mixed) == len(mixer_coefficients) - 1:
                mixed.append(sample_size - added)
                break
            k = floor(sample_size * coeff)
            added += k
            mixed.append(k)
        assert sum(mixed) == sample_size
        return mixed

    def forward(self, x, y, model, sample_size, batch_size, mixer_coefficients, variates=None):
        assert len(mixer_coefficients) == len(self.samplers), '[ERROR] Each sampler must have a mixer coefficient'
        assert abs(sum(mixer_coefficients) - 1.0) < 1e-8, '[ERROR] mixer coefficeints need to sum to one.'
        counts = self._coefficient_to_num_samples(sample_size, mixer_coefficients)
        all_samples = []
        for h_dim in self.h_dims:
            h = torch.empty(batch_size, sample_size, h_dim, dtype=torch.float32, device=x.device)
            s0 = 0
            for i, (sampler, n) in enumerate(zip(self.samplers, counts)):
                if n == 0:
                    continue
                e = variates[i].to(x.device) if variates is not None else sampler.draw(batch_size, n, h_dim, x.device)
                ops.sampler_fill(sampler.kind, e, y, h, s0)
                s0 += n
            all_samples.append(h.flatten(0, 1))
        return torch.Tensor(), all_samples
