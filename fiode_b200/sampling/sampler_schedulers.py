"""sampling/sampler_schedulers.py:14-77 of the reference -- host-side scalars, kept on the host."""
import numpy as np


class LinearScheduler:
    def __init__(self, rate, bias=0., clamp='min', clamp_val=0., start=0):
        assert clamp_val >= 0, 'Schedulers must return positive number'
        self.rate, self.bias, self.clamp, self.clamp_val, self.start = rate, bias, clamp, clamp_val, start

    def sampler_weight(self, epoch_num):
        if epoch_num < self.start:
            return 0. if self.rate > 0 else 1.
        weight = (epoch_num - self.start) * self.rate + self.bias
        if self.clamp == 'max':
            return min(weight, self.clamp_val)
        if self.clamp == 'min':
            return max(weight, self.clamp_val)
        return weight


class ConstantScheduler:
    def __init__(self, constant):
        assert constant >= 0, 'Schedulers must return positive number'
        self.constant = constant

    def sampler_weight(self, epoch_num):
        return self.constant


class SwitchScheduler:
    def __init__(self, start, end, trigger):
        assert start >= 0 and end >= 0, 'Schedulers must return positive number'
        self.start, self.end, self.trigger = start, end, trigger

    def sampler_weight(self, epoch_num):
        return self.start if epoch_num < self.trigger else self.end


class CompositeSamplerScheduler:
    def __init__(self, schedulers, scheduler_weights):
        assert len(schedulers) == len(scheduler_weights), 'each scheduler needs a weight'
        self.schedulers = schedulers
        self.scheduler_weights = np.array(scheduler_weights)

    def get_mixer_coefficients(self, epoch_num):
        c = np.array([s.sampler_weight(epoch_num) for s in self.schedulers]) * self.scheduler_weights
        return c / (np.linalg.norm(c, ord=1) + 1e-12)
