"""ctypes binding of libfiode_b200.so (the C ABI declared in include/fiode_b200.h).

There is NO fallback: if the shared library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / 'libfiode_b200.so'

_lib = None

P, I, LL, F32, F64, SZ = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_double, C.c_size_t


class FiodeError(RuntimeError):
    pass


class DynWeights(C.Structure):
    _fields_ = [('Q1', P), ('b1', P), ('Qx', P), ('bx', P), ('Q2', P), ('b2', P), ('Q3', P), ('b3', P),
                ('alpha_1', F32), ('alpha_2', F32), ('sigma_1', F32),
                ('scale_nominal', I), ('max_iter', I), ('tol', F32)]


class FiOptions(C.Structure):
    _fields_ = [('kappa_dev', P), ('dropout_p', F32), ('seed', C.c_ulonglong), ('step_dev', P)]


class BackboneWeights(C.Structure):
    _fields_ = [(n, P) for n in ('mu', 'std', 'Qf0', 'b0', 'Q2', 'b2', 'Qf4', 'b4', 'Q6', 'b6',
                                 'Q9', 'b9', 'Q11', 'b11', 'Qfc', 'bfc', 'packed')]


class SolveStats(C.Structure):
    _fields_ = [('n_accept', I), ('n_reject', I), ('nfe', I), ('status', I),
                ('t_final', F64), ('last_dt', F64), ('n_logged', I), ('pad_', I),
                ('dts', F64 * 64), ('ratios', F32 * 64)]


class SegwayWeights(C.Structure):
    _fields_ = [('W0', P), ('b0', P), ('W1', P), ('b1', P), ('W2', P), ('b2', P), ('P', P)]


# name -> (restype, argtypes): every entry point of include/fiode_b200.h
SIGNATURES = {
    'fiode_last_error': (C.c_char_p, []),
    'fiode_version': (I, []),
    'fiode_launch_count': (LL, []),
    'fiode_reset_launch_count': (None, []),
    'fiode_set_side_streams': (None, [I]),
    'fiode_cayley_linear_workspace_bytes': (SZ, [I, I]),
    'fiode_cayley_linear': (I, [P, P, I, I, P, P, SZ, P]),
    'fiode_conv_alpha_init': (I, [P, I, I, I, I, P, P, SZ, P]),
    'fiode_cayley_conv_workspace_bytes': (SZ, [I, I, I, I]),
    'fiode_cayley_conv': (I, [P, P, I, I, I, I, P, P, SZ, P]),
    'fiode_backbone_workspace_bytes': (SZ, [I]),
    'fiode_backbone_slab_images': (I, []),
    'fiode_backbone_set_conv0_path': (None, [I]),
    'fiode_debug_conv0': (I, [C.POINTER(BackboneWeights), P, I, I, P, P, SZ, P]),
    'fiode_backbone_packed_bytes': (SZ, []),
    'fiode_backbone_pack': (I, [C.POINTER(BackboneWeights), P, SZ, P]),
    'fiode_backbone_fwd': (I, [C.POINTER(BackboneWeights), P, I, P, P, SZ, P]),
    'fiode_backbone_fwd_u8': (I, [C.POINTER(BackboneWeights), P, I, P, P, SZ, P]),
    'fiode_backbone_enable_stage_timing': (None, [I]),
    'fiode_backbone_stage_times': (I, [P, I]),
    'fiode_cayley_conv_fwd_workspace_bytes': (SZ, [I, I, I, I]),
    'fiode_cayley_conv_fwd': (I, [P, P, P, I, I, I, I, I, P, P, SZ, P]),
    'fiode_linear_workspace_bytes': (SZ, [I, I, I]),
    'fiode_linear_fwd': (I, [P, P, P, I, I, I, I, P, P, SZ, P]),
    'fiode_linear_static_max_bytes': (SZ, []),
    'fiode_linear_fwd_static': (I, [P, P, P, I, I, I, I, P, P]),
    'fiode_conv1x1_s2d_fwd': (I, [P, P, P, I, I, I, I, I, P, P]),
    'fiode_space_to_depth': (I, [P, I, I, I, P, P]),
    'fiode_group_sort': (I, [P, I, I, I, P, P]),
    'fiode_normalize': (I, [P, I, I, I, P, P, P, P]),
    'fiode_group_sort_bwd': (I, [P, P, I, I, I, P, P]),
    'fiode_bias_grad': (I, [P, I, I, I, P, P]),
    'fiode_linear_bwd_workspace_bytes': (SZ, [I, I, I]),
    'fiode_linear_bwd': (I, [P, P, P, I, I, I, P, P, P, P, SZ, P]),
    'fiode_conv1x1_s2d_bwd_workspace_bytes': (SZ, [I, I, I, I]),
    'fiode_conv1x1_s2d_bwd': (I, [P, P, P, I, I, I, I, P, P, P, P, SZ, P]),
    'fiode_cayley_conv_layer_bwd_workspace_bytes': (SZ, [I, I, I, I]),
    'fiode_cayley_conv_layer_bwd': (I, [P, P, P, I, I, I, I, P, P, P, P, SZ, P]),
    'fiode_cayley_linear_bwd_workspace_bytes': (SZ, [I, I]),
    'fiode_cayley_linear_bwd': (I, [P, P, I, I, P, P, P, P, SZ, P]),
    'fiode_cayley_linear_save': (I, [P, P, I, I, P, P, P, SZ, P]),
    'fiode_cayley_linear_bwd_saved': (I, [P, P, I, I, P, P, P, P, P, SZ, P]),
    'fiode_cayley_linear_ctx_floats': (SZ, [I, I]),
    'fiode_cayley_linear_fwd_ctx': (I, [P, P, I, I, P, P, P, SZ, P]),
    'fiode_cayley_linear_bwd_ctx': (I, [P, P, I, I, P, P, P, P, P, SZ, P]),
    'fiode_cayley_small_ok': (I, [I, I]),
    'fiode_cayley_small_batch_fwd': (I, [I, P, P, P, P, P, P, P, P]),
    'fiode_cayley_small_batch_bwd': (I, [I, P, P, P, P, P, P, P, P, P, P]),
    'fiode_cayley_conv_bwd_workspace_bytes': (SZ, [I, I, I, I]),
    'fiode_cayley_conv_bwd': (I, [P, P, I, I, I, I, P, P, P, P, SZ, P]),
    'fiode_rhs_workspace_bytes': (SZ, [LL]),
    'fiode_rhs_fwd': (I, [C.POINTER(DynWeights), P, P, LL, LL, LL, P, P, P, P, P, SZ, P]),
    'fiode_barrier_project_fwd': (I, [P, P, P, LL, I, F32, P, P, P, P, SZ, P]),
    'fiode_barrier_project_bwd': (I, [P, P, P, P, LL, P, P, P]),
    'fiode_barrier_project_bwd_upper': (I, [P, P, P, P, LL, P, P, P, P]),
    'fiode_certify_workspace_bytes': (SZ, [I, I]),
    'fiode_set_mlp_path': (None, [I]),
    'fiode_get_mlp_path': (I, []),
    'fiode_certify_grid': (I, [C.POINTER(DynWeights), P, LL, I, P, P, P, I, I, F32, P, P, P, SZ, P]),
    'fiode_sampler_fill': (I, [I, P, P, I, I, I, I, P, P]),
    'fiode_decision_boundary': (I, [P, P, LL, I, I, P, P]),
    'fiode_fi_loss_workspace_bytes': (SZ, [LL, I]),
    'fiode_fi_loss_fwd': (I, [C.POINTER(DynWeights), P, P, P, I, I, F32, P, P, P, P, P, P, P, P, P, P, P, SZ, P]),
    'fiode_fi_loss_fwd_opt': (I, [C.POINTER(DynWeights), P, P, P, I, I, F32, P, P, C.POINTER(FiOptions), P, P, P, P, P, P, P, P, P, SZ, P]),
    'fiode_fi_loss_bwd_opt': (I, [C.POINTER(DynWeights), P, P, P, I, I, F32, F32, P, P, C.POINTER(FiOptions), P, P, P, P, P, P, P, P, P, P, P, P, P, P, P,
                                  P, SZ, P]),
    'fiode_dropout_mask': (I, [F32, C.c_ulonglong, P, I, LL, P, P]),
    'fiode_fi_loss_bwd': (I, [C.POINTER(DynWeights), P, P, P, I, I, F32, F32, P, P, P, P, P, P, P, P, P,
                              P, P, P, P, P, P, P, P, P, SZ, P]),
    'fiode_segway_grid_workspace_bytes': (SZ, []),
    'fiode_segway_grid': (I, [C.POINTER(SegwayWeights), P, I, P, I, P, I, F32, F32, LL, LL, LL, P, P, P, P, SZ, P]),
    'fiode_segway_simulate': (I, [C.POINTER(SegwayWeights), P, I, P, I, F64, F64, P, P, P, P]),
    'fiode_segway_param_count': (I, []),
    'fiode_segway_hdot_bwd_workspace_bytes': (SZ, [LL]),
    'fiode_segway_hdot': (I, [P, P, LL, P, P]),
    'fiode_segway_hdot_bwd': (I, [P, P, LL, P, P, P, P, SZ, P]),
    'fiode_segway_pgd_step': (I, [P, P, P, LL, F32, F32, F32, F32, P]),
    'fiode_grid_count': (LL, [I, I]),
    'fiode_grid_workspace_bytes': (SZ, [I, I]),
    'fiode_grid_enumerate': (I, [I, I, LL, LL, P, P, SZ, P]),
    'fiode_dopri5_workspace_bytes': (SZ, [I]),
    'fiode_dopri5_simplex': (I, [C.POINTER(DynWeights), P, P, I, I, P, I, F64, F64, LL, P, P, P, SZ, I, P]),
    'fiode_dopri5_simplex_logged': (I, [C.POINTER(DynWeights), P, P, I, I, P, I, F64, F64, LL, P, P, P, SZ, I, P, P, LL, P]),
    'fiode_dopri5_max_concurrent_solves': (I, [I, I]),
    'fiode_rhs_train_workspace_bytes': (SZ, [LL]),
    'fiode_rhs_fwd_train': (I, [C.POINTER(DynWeights), P, P, LL, P, P, P, P, P, P, SZ, P]),
    'fiode_dyn_transposes': (I, [C.POINTER(DynWeights), P, P, P, P, P]),
    'fiode_rhs_bwd_data': (I, [C.POINTER(DynWeights), P, LL, P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, SZ, P]),
    'fiode_rhs_bwd_params_workspace_bytes': (SZ, [LL]),
    'fiode_rhs_bwd_params': (I, [P, P, P, P, P, P, P, LL, P, P, P, P, P, P, P, P, SZ, P]),
    'fiode_lincomb': (I, [P, LL, I, P, P, P]),
    'fiode_ibp_workspace_bytes': (SZ, []),
    'fiode_ibp_sigmoid': (I, [P, P, P, LL, F32, F32, F32, F32, P, P, P]),
    'fiode_ibp_cbf_qp': (I, [P, P, P, LL, F32, F32, F32, I, F32, P, P, P, P, SZ, P]),
    'fiode_runner_up_mask': (I, [P, LL, I, F32, P, P]),
    'fiode_perturbed_vdot': (I, [P, P, P, I, P, LL, P, P, P, SZ, P]),
    'fiode_dopri5_simplex_batched': (I, [C.POINTER(DynWeights), P, P, I, I, I, P, I, F64, F64, LL, P, P, I, P]),
}


def declared_symbols():
    """Entry points named in include/fiode_b200.h (parsed, so the header stays the source of truth)."""
    import re
    hdr = (_HERE.parent / 'include' / 'fiode_b200.h').read_text()
    return sorted(set(re.findall(r'\b(fiode_[a-z0-9_]+)\s*\(', hdr)))


def lib() -> C.CDLL:
    """Load the library (once).  Raises FiodeError if it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise FiodeError(f'{LIB_PATH} is missing: run `python -m fiode_b200.build` '
                             '(there is no CPU / PyTorch fallback for the fiode_b200 hot path)')
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int, what: str = ''):
    if rc != 0:
        msg = lib().fiode_last_error().decode(errors='replace')
        raise FiodeError(f'{what} failed (rc={rc}): {msg}')


def ptr(t):
    """Raw device pointer of a contiguous CUDA tensor (or None)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise FiodeError('fiode_b200 runs on CUDA tensors only (no CPU fallback); got a CPU tensor')
    assert t.is_contiguous(), 'fiode_b200 needs contiguous tensors'
    return C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count() -> int:
    return int(lib().fiode_launch_count())


def reset_launch_count():
    lib().fiode_reset_launch_count()
