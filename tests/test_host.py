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


def test_segway_training_mirrors_have_reference_keys():
    """NNController / LyaQuadratic mirrors (fiode_b200/control/segway.py) carry the reference's state_dict keys, so the files
    train_segway.py saves and certify_segway.py loads are interchangeable."""
    from fiode_b200.control import segway
    from tests import fixtures as fx
    g = fx.load_golden('segway_train.pt')
    ctrl = segway.NNController(None, 3, 1, 32)
    assert set(ctrl.state_dict().keys()) == set(g['ctrl'].keys())
    ctrl.load_state_dict(g['ctrl'])
    lya = segway.LyaQuadratic(g['P'])
    assert list(lya.state_dict().keys()) == ['P.weight'] and torch.equal(lya.P.weight, g['P'])
    with pytest.raises(AssertionError):
        segway.NNController(None, 4, 1, 32)


def test_sweep_verdict_logic():
    """tools/sweep.py: equal / noise / different against an oracle row and its 1e-6-noise band."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location('sweep_tool', os.path.join(os.path.dirname(os.path.dirname(__file__)), 'tools', 'sweep.py'))
    # the module imports bench + fiode_b200.ops at import time (no GPU needed for that)
    sw = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sw)
    row = dict(oracle=[52, 0, 314], oracle_noise_1e6=[[52, 0, 314], [53, 1, 326]])
    assert sw.verdict((52, 0, 314), row) == 'equal'
    assert sw.verdict((53, 1, 326), row) == 'noise'
    assert sw.verdict((54, 2, 338), row) == 'noise'          # band (1 step) doubled plus one step
    assert sw.verdict((60, 0, 362), row) == 'different'
    assert sw.verdict((52, 0, 314), None) == ''
    rows = [dict(case='README', launch='cooperative', B=64, tol=1e-3, n_accept=3, n_reject=0, nfe=20, oracle=[3, 0, 20], verdict='equal',
                 rows_per_s=1.0, tflops=0.1, hbm_gbs=0.1),
            dict(case='README', launch='cooperative', B=64, tol=1e-2, status=1)]
    s = sw.summary(rows)
    assert s['equal'] == 1 and s['failed'] == 1 and s['cells'] == 2 and s['worst']['gpu'] == [3, 0, 20]
