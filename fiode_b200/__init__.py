"""fiode_b200 -- B200-native (sm_100a) implementation of FI-ODE's data-parallel hot path.

The module tree mirrors the reference's names so that Hydra ``_target_`` paths can be retargeted
(SURVEY.md section 8b):

    reference                                              here
    libs.ortho_conv.layers.{cayley,CayleyConv,...}      -> fiode_b200.layers
    libs.ortho_conv.models_ortho.KWLarge_Concat, models -> fiode_b200.models
    dynamics.classification.OrthoClassDynProjectSimplexLips -> fiode_b200.dynamics.classification
    dynamics.init_coordinates / output_coordinates      -> fiode_b200.dynamics.*
    barrier_projection.barrier_projection               -> fiode_b200.barrier_projection
    torchdiffeq.odeint                                  -> fiode_b200.odeint.odeint

All computation happens in libfiode_b200.so (hand-written CUDA behind the C ABI of
include/fiode_b200.h); there is no CPU or eager-PyTorch fallback.
"""
from ._lib import FiodeError, lib, launch_count, reset_launch_count  # noqa: F401

__all__ = ['FiodeError', 'lib', 'launch_count', 'reset_launch_count']
