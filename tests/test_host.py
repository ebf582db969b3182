"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol that
include/fiode_b200.h declares, the ctypes table covers them, host-only entry points work, and the
product path fails loudly (no CPU fallback)."""
import pytest
import torch

from tests import fixtures as fx


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    ge.build()
    from fiode_b200 import _lib
    L = _lib.lib()
    declared = _lib.declared_symbols()
    assert len(declared) >= 29
    for name in declared:
        assert hasattr(L, name), name
        assert name in _lib.SIGNATURES, f'{name} has no ctypes signature'
    assert L.fiode_version() >= 100
    assert L.fiode_last_error() is not None


def test_grid_count_matches_reference_counter():
    """fiode_grid_count is host-only; the golden counts come from the reference's own
    count_samples_decision_boundary (robustness/eval_utils.py:72-89)."""
    from fiode_b200 import grid
    g = fx.load_golden('certify.pt')
    for (n, T), want in g['counts'].items():
        assert grid.count_samples_decision_boundary(n, T) == want, (n, T)
    assert grid.count_samples_decision_boundary(10, 40) == 41320837


def test_no_cpu_fallback():
    from fiode_b200 import ops, FiodeError
    from fiode_b200.layers import CayleyLinear
    lin = CayleyLinear(10, 128)
    with pytest.raises(FiodeError, match='no CPU fallback'):
        lin(torch.zeros(4, 10))
    with pytest.raises(FiodeError):
        ops.barrier_project_fwd(torch.zeros(4, 10), torch.zeros(4, 10))


def test_error_string_convention():
    from fiode_b200 import _lib
    L = _lib.lib()
    rc = L.fiode_grid_count(1, 40)          # n < 2 is rejected on the host, before any CUDA call
    assert rc == -1 and b'fiode_grid_count' in L.fiode_last_error()


def test_state_dict_keys_match_reference_layout(sd):
    """The mirrored module tree loads a reference-layout (Lightning) state_dict: same names, same shapes."""
    ivp = fx.build_b200_model(sd, device='cpu')
    own = ivp.state_dict()
    for k, v in sd.items():
        k2 = k[len('model.'):]
        assert k2 in own and tuple(own[k2].shape) == tuple(v.shape), k
        assert torch.equal(own[k2], v), k


def test_grid_pt_reader_round_trip(tmp_path):
    """a reference-format grid file (list of 10 float32 (N,10) tensors, label i = columns 0 and i swapped,
    sample_decision_boundary.py:37-42) read back into the canonical uint8 numerators; built here from the oracle grid."""
    from fiode_b200 import grid
    from oracle import certify as oc
    T = 10
    z = oc.boundary_grid_numerators(10, T)
    grids = [oc.grid_for_label(z, T, i).float() for i in range(10)]
    f = tmp_path / 'grid_10.pt'
    torch.save(grids, f)
    for label in (0, 3, 9):
        back = grid.numerators_from_grid_pt(str(f), T, label=label, device='cpu')
        assert back.dtype == torch.uint8 and torch.equal(back.long(), torch.as_tensor(z).long())
    bad = [g.clone() for g in grids]
    bad[0][5, 0] += 0.1
    with pytest.raises(ValueError):
        grid.numerators_from_grid_pt(bad, T, device='cpu')


def test_lightning_checkpoint_loads(sd, tmp_path):
    """a Lightning-style model.ckpt ({'state_dict': {'model.*': ...}, plus trainer keys and tensors the hot path does
    not own}) loads into the mirrored IVP and into LyapunovLearning with strict=False, like eval_utils.py:92-107."""
    from fiode_b200 import checkpoint, pl_modules, lya_cands
    from fiode_b200.sampling import sampler as smp, sampler_schedulers as sch
    from fiode_b200.dynamics.output_coordinates import DefaultOutputFun
    extra = {'some_callback.buffer': torch.zeros(3)}
    f = tmp_path / 'model.ckpt'
    torch.save({'epoch': 7, 'global_step': 1234, 'state_dict': {**sd, **extra}, 'optimizer_states': []}, f)
    ivp = fx.build_b200_model({k: torch.zeros_like(v) for k, v in sd.items()}, device='cpu')
    missing, unexpected = checkpoint.load_reference_checkpoint(ivp, str(f))
    assert all('static_state' in m for m in missing) and not unexpected
    for k, v in sd.items():
        assert torch.equal(ivp.state_dict()[k[len('model.'):]], v), k
    ivp2 = fx.build_b200_model({k: torch.zeros_like(v) for k, v in sd.items()}, device='cpu')
    mod = pl_modules.LyapunovLearning(
        order=1, h_sample_size=8, h_dist_lim=15, lya_cand=lya_cands.DecisionBoundary(on_simplex=True),
        sampler=smp.CompositeSampler((10,), [smp.UniformSimplexSampling((10,))]),
        sampler_scheduler=sch.CompositeSamplerScheduler([sch.ConstantScheduler(1.0)], [1.]),
        dynamics=ivp2.dyn_fun, output=DefaultOutputFun(), init_fun=ivp2.init_coordinates, t_max=1.0)
    missing, unexpected = checkpoint.load_reference_checkpoint(mod, str(f))
    assert unexpected == ['some_callback.buffer']
    assert torch.equal(mod.model.dyn_fun.mlp_to_mlp.weight, sd['model.dyn_fun.mlp_to_mlp.weight'])
