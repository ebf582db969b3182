"""fiode_linear_fwd_static: the F.linear product with its scratch in a built-in device array (no workspace argument, no
allocation inside the call) against fiode_linear_fwd with a caller workspace, through the C ABI.  No counterpart in the
reference (added for callers that cannot allocate).  Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = 'cuda'


def _randn(*shape, seed):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)).to(DEV)


def _both(M, N, K, gs, seed):
    """(y from fiode_linear_fwd, y from fiode_linear_fwd_static) on the same seeded inputs, straight through ctypes."""
    from fiode_b200._lib import check, lib, ptr, stream_ptr
    L = lib()
    X, Q, b = _randn(M, K, seed=seed), _randn(N, K, seed=seed + 1) / K ** 0.5, _randn(N, seed=seed + 2)
    nb = L.fiode_linear_workspace_bytes(M, N, K)
    ws = torch.empty(nb, dtype=torch.uint8, device=DEV)
    want = torch.full((M, N), float('nan'), device=DEV)
    got = torch.full((M, N), float('nan'), device=DEV)
    check(L.fiode_linear_fwd(ptr(X), ptr(Q), ptr(b), M, N, K, gs, ptr(want), ptr(ws), nb, stream_ptr()), 'fiode_linear_fwd')
    torch.cuda.synchronize()
    stats0 = torch.cuda.memory_stats().get('allocation.all.allocated', 0)
    check(L.fiode_linear_fwd_static(ptr(X), ptr(Q), ptr(b), M, N, K, gs, ptr(got), stream_ptr()), 'fiode_linear_fwd_static')
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats().get('allocation.all.allocated', 0) == stats0, 'the static entry point went through torch'
    return want, got


# split-K FFMA product with GroupSort (8 MiB of partials), the tensor-core product (re-tiled weights in the scratch), the fc head
@pytest.mark.parametrize('M,N,K,gs', [(128, 512, 4096, 1), (2048, 512, 512, 1), (64, 10, 512, 0)])
def test_linear_fwd_static_equals_linear_fwd_twice_in_a_row(M, N, K, gs):
    from fiode_b200._lib import lib
    assert lib().fiode_linear_static_max_bytes() == 16 << 20
    for call in range(2):              # the second call finds the first one's contents in the shared scratch
        want, got = _both(M, N, K, gs, seed=100 + 10 * call)
        assert torch.isfinite(got).all()
        assert torch.equal(got, want), (call, float((got - want).abs().max()))


def test_linear_fwd_static_allocates_nothing_and_rejects_oversize():
    from fiode_b200._lib import lib, ptr, stream_ptr
    L = lib()
    M, N, K = 256, 512, 512
    _both(M, N, K, 1, seed=7)                                   # kernels loaded, scratch address resolved
    X, Q, y = _randn(M, K, seed=1), _randn(N, K, seed=2), torch.empty(M, N, device=DEV)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    assert L.fiode_linear_fwd_static(ptr(X), ptr(Q), None, M, N, K, 1, ptr(y), stream_ptr()) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, 'device memory changed hands during the call'
    # beyond the built-in array: a clear error, no launch
    M, N, K = 512, 512, 4096            # 16 split-K partials of 512 x 512 floats + 256 bytes
    assert L.fiode_linear_workspace_bytes(M, N, K) > L.fiode_linear_static_max_bytes()
    X, Q, y = _randn(M, K, seed=3), _randn(N, K, seed=4), torch.zeros(M, N, device=DEV)
    assert L.fiode_linear_fwd_static(ptr(X), ptr(Q), None, M, N, K, 0, ptr(y), stream_ptr()) != 0
    msg = L.fiode_last_error().decode()
    assert 'fiode_linear_fwd_static' in msg and 'built-in array' in msg, msg
    torch.cuda.synchronize()
    assert not y.any()
