"""Tensor-level wrappers over the C ABI (include/fiode_b200.h).  PyTorch is used for device memory
and streams only; every computation below happens in libfiode_b200.so.  CUDA tensors only."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import DynWeights, BackboneWeights, SolveStats, check, lib, ptr, stream_ptr

NH = 10


def _f32(t):
    return t.detach().to(torch.float32).contiguous()


def _ws(nbytes, device):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


# ---- Cayley weight preparation ------------------------------------------------------------------
def cayley_linear(weight, alpha, save_minv=False, save_ctx=False):
    """Q = cayley(alpha * W / ||W||)  (libs/ortho_conv/layers.py:103-104).  ``save_minv``: also return (I + A)^-1 for the
    adjoint (``cayley_linear_bwd(..., minv=)``), as autograd keeps the forward's intermediates; ``save_ctx``: return the
    library's whole context buffer instead (scale, inverse, staged matrices: ``cayley_linear_bwd(..., ctx=)``)."""
    W = _f32(weight)
    cout, cin = W.shape
    a = _f32(alpha).reshape(1)
    Q = torch.empty_like(W)
    n = lib().fiode_cayley_linear_workspace_bytes(cout, cin)
    ws = _ws(n, W.device)
    if save_ctx:
        ctx = torch.empty(lib().fiode_cayley_linear_ctx_floats(cout, cin), dtype=torch.float32, device=W.device)
        check(lib().fiode_cayley_linear_fwd_ctx(ptr(W), ptr(a), cout, cin, ptr(Q), ptr(ctx), ptr(ws), n, stream_ptr()),
              'fiode_cayley_linear_fwd_ctx')
        return Q, ctx
    if save_minv:
        m = min(cout, cin)
        minv = torch.empty(m, m, dtype=torch.float32, device=W.device)
        check(lib().fiode_cayley_linear_save(ptr(W), ptr(a), cout, cin, ptr(Q), ptr(minv), ptr(ws), n, stream_ptr()),
              'fiode_cayley_linear_save')
        return Q, minv
    check(lib().fiode_cayley_linear(ptr(W), ptr(a), cout, cin, ptr(Q), ptr(ws), n, stream_ptr()), 'fiode_cayley_linear')
    return Q


def conv_alpha_init(weight, n):
    """CayleyConv's lazily created alpha = ||wfft||_F (layers.py:80-81) as a 0-dim device tensor."""
    w = _f32(weight)
    cout, cin, k, _ = w.shape
    out = torch.empty(1, dtype=torch.float32, device=w.device)
    nb = lib().fiode_cayley_conv_workspace_bytes(cout, cin, max(k, 2), n)
    ws = _ws(nb, w.device)
    check(lib().fiode_conv_alpha_init(ptr(w), cout, cin, k, n, ptr(out), ptr(ws), nb, stream_ptr()),
          'fiode_conv_alpha_init')
    return out.reshape(())


def cayley_conv(weight, alpha, n):
    """Per-frequency orthogonal matrices of a CayleyConv (layers.py:74-82).  k=3 -> complex64
    (F, cout, cin); k=1 -> real (cout, cin) shared by every frequency."""
    w = _f32(weight)
    cout, cin, k, _ = w.shape
    a = _f32(alpha).reshape(1)
    F = n * (n // 2 + 1)
    if k == 1:
        Q = torch.empty(cout, cin, dtype=torch.float32, device=w.device)
    else:
        Q = torch.empty(F, cout, cin, 2, dtype=torch.float32, device=w.device)
    nb = lib().fiode_cayley_conv_workspace_bytes(cout, cin, k, n)
    ws = _ws(nb, w.device)
    check(lib().fiode_cayley_conv(ptr(w), ptr(a), cout, cin, k, n, ptr(Q), ptr(ws), nb, stream_ptr()), 'fiode_cayley_conv')
    return Q if k == 1 else torch.view_as_complex(Q)


# ---- layers ---------------------------------------------------------------------------------------
def cayley_conv_fwd(x, Qf, bias, group_sort=False):
    x = _f32(x)
    B, cin, n, _ = x.shape
    Qr = torch.view_as_real(Qf).contiguous() if Qf.is_complex() else Qf.contiguous()
    cout = Qr.shape[1]
    y = torch.empty(B, cout, n, n, dtype=torch.float32, device=x.device)
    nb = lib().fiode_cayley_conv_fwd_workspace_bytes(B, cin, cout, n)
    ws = _ws(nb, x.device)
    check(lib().fiode_cayley_conv_fwd(ptr(x), ptr(Qr), ptr(_f32(bias)) if bias is not None else None, B, cin, cout, n,
                                      int(group_sort), ptr(y), ptr(ws), nb, stream_ptr()), 'fiode_cayley_conv_fwd')
    return y


def conv1x1_s2d_fwd(x, Q, bias, group_sort=False):
    x = _f32(x)
    B, cin, n, _ = x.shape
    cout = Q.shape[0]
    y = torch.empty(B, cout, n // 2, n // 2, dtype=torch.float32, device=x.device)
    check(lib().fiode_conv1x1_s2d_fwd(ptr(x), ptr(_f32(Q)), ptr(_f32(bias)) if bias is not None else None, B, cin, cout,
                                      n, int(group_sort), ptr(y), stream_ptr()), 'fiode_conv1x1_s2d_fwd')
    return y


def linear_fwd(X, Q, bias, group_sort=False):
    X = _f32(X)
    lead = X.shape[:-1]
    X2 = X.reshape(-1, X.shape[-1])
    M, K = X2.shape
    N = Q.shape[0]
    y = torch.empty(M, N, dtype=torch.float32, device=X.device)
    nb = lib().fiode_linear_workspace_bytes(M, N, K)
    ws = _ws(nb, X.device)
    check(lib().fiode_linear_fwd(ptr(X2), ptr(_f32(Q)), ptr(_f32(bias)) if bias is not None else None, M, N, K,
                                 int(group_sort), ptr(y), ptr(ws), nb, stream_ptr()), 'fiode_linear_fwd')
    return y.reshape(*lead, N)


def space_to_depth(x):
    x = _f32(x)
    B, Cc, n, _ = x.shape
    y = torch.empty(B, 4 * Cc, n // 2, n // 2, dtype=torch.float32, device=x.device)
    check(lib().fiode_space_to_depth(ptr(x), B, Cc, n, ptr(y), stream_ptr()), 'fiode_space_to_depth')
    return y


def group_sort(x):
    x = _f32(x)
    B, Cc = x.shape[:2]
    inner = x[0, 0].numel() if x.dim() > 2 else 1
    y = torch.empty_like(x)
    check(lib().fiode_group_sort(ptr(x), B, Cc, inner, ptr(y), stream_ptr()), 'fiode_group_sort')
    return y


def backbone_pack(bw: BackboneWeights, device):
    """Kernel-ready copies of the prepared backbone matrices (fiode_backbone_pack); pass the block as ``bw.packed``."""
    nb = lib().fiode_backbone_packed_bytes()
    blk = torch.empty(nb, dtype=torch.uint8, device=device)
    check(lib().fiode_backbone_pack(C.byref(bw), ptr(blk), nb, stream_ptr()), 'fiode_backbone_pack')
    return blk


def backbone_fwd(bw: BackboneWeights, x, keep=()):
    """s = param_map(x) for the CIFAR KWLarge_Concat backbone.  ``keep`` holds the tensors the struct
    points into (so they outlive the asynchronous launch).  x: float32 in [0, 1], or the dataset's uint8 pixels
    (ToTensor's / 255 then happens inside the first kernel; same bits as the float call on x / 255)."""
    u8 = x.dtype == torch.uint8
    x = x.contiguous() if u8 else _f32(x)
    B = x.shape[0]
    s = torch.empty(B, NH, dtype=torch.float32, device=x.device)
    nb = lib().fiode_backbone_workspace_bytes(B)
    ws = _ws(nb, x.device)
    fn = lib().fiode_backbone_fwd_u8 if u8 else lib().fiode_backbone_fwd
    check(fn(C.byref(bw), ptr(x), B, ptr(s), ptr(ws), nb, stream_ptr()), 'fiode_backbone_fwd')
    return s


# ---- dynamics -------------------------------------------------------------------------------------
def rhs_fwd(dw: DynWeights, h, s, s_group=None, return_aux=False):
    """f = eval_dot_light(h, s) (dynamics/classification.py:117-126).  s: (1,10) broadcast, (N,10)
    per row, or (G,10) with ``s_group`` consecutive rows of h sharing one s row."""
    h = _f32(h)
    s = _f32(s)
    N = h.shape[0]
    if s.shape[0] == 1:
        s_rows, sg = 1, N
    elif s_group is not None:
        s_rows, sg = s.shape[0], int(s_group)
    else:
        assert s.shape[0] == N, 'static state must have 1 or N rows'
        s_rows, sg = N, 1
    f = torch.empty_like(h)
    ft = torch.empty_like(h) if return_aux else None
    mu = torch.empty(N, dtype=torch.float32, device=h.device) if return_aux else None
    it = torch.empty(1, dtype=torch.int32, device=h.device) if return_aux else None
    nb = lib().fiode_rhs_workspace_bytes(N)
    ws = _ws(nb, h.device)
    check(lib().fiode_rhs_fwd(C.byref(dw), ptr(h), ptr(s), N, s_rows, sg, ptr(f), ptr(ft), ptr(mu), ptr(it), ptr(ws), nb,
                              stream_ptr()), 'fiode_rhs_fwd')
    return (f, ft, mu, it) if return_aux else f


def barrier_project_fwd(lower, nominal, upper=None, max_iter=30, tol=1e-4):
    lower, nominal = _f32(lower), _f32(nominal)
    upper = _f32(upper) if upper is not None else None
    N = nominal.shape[0]
    v = torch.empty_like(nominal)
    mu = torch.empty(N, dtype=torch.float32, device=nominal.device)
    it = torch.empty(1, dtype=torch.int32, device=nominal.device)
    ws = _ws(256, nominal.device)
    check(lib().fiode_barrier_project_fwd(ptr(lower), ptr(upper), ptr(nominal), N, max_iter, tol, ptr(v), ptr(mu), ptr(it),
                                          ptr(ws), 256, stream_ptr()), 'fiode_barrier_project_fwd')
    return v, mu, it


def barrier_project_bwd(grad_v, v, mu, nominal):
    grad_v, v, mu, nominal = _f32(grad_v), _f32(v), _f32(mu), _f32(nominal)
    N = nominal.shape[0]
    dl, dn = torch.empty_like(nominal), torch.empty_like(nominal)
    check(lib().fiode_barrier_project_bwd(ptr(grad_v), ptr(v), ptr(mu), ptr(nominal), N, ptr(dl), ptr(dn), stream_ptr()),
          'fiode_barrier_project_bwd')
    return dl, dn


def barrier_project_bwd_upper(grad_v, v, mu, nominal):
    """(d_lower, d_upper, d_nominal) of FastBarrierProjection (barrier_projection.py:164-214) from its saved forward state."""
    grad_v, v, mu, nominal = _f32(grad_v), _f32(v), _f32(mu), _f32(nominal)
    dl, du, dn = torch.empty_like(v), torch.empty_like(v), torch.empty_like(v)
    check(lib().fiode_barrier_project_bwd_upper(ptr(grad_v), ptr(v), ptr(mu), ptr(nominal), v.shape[0], ptr(dl), ptr(du), ptr(dn),
                                                stream_ptr()), 'fiode_barrier_project_bwd_upper')
    return dl, du, dn


def certify_grid(dw: DynWeights, gridz, T, s, labels, kappa, batches=10, return_iters=False):
    """Per-image maxima of the Lipschitz grid check (robustness/certify_lipschitz.py:104-143).
    gridz: (Ng,10) uint8 numerators of the canonical label-0 grid; returns (M,2)."""
    assert gridz.dtype == torch.uint8 and gridz.is_cuda
    gridz = gridz.contiguous()
    s = _f32(s)
    labels = labels.to(torch.int32).contiguous()
    M, Ng = s.shape[0], gridz.shape[0]
    lut = (torch.arange(T + 1, dtype=torch.float64) / T).float().to(s.device)   # what grid.float() holds
    n_chunks = batches + (1 if Ng % batches else 0)
    out = torch.empty(M, 2, dtype=torch.float32, device=s.device)
    iters = torch.empty(M, n_chunks, dtype=torch.int32, device=s.device) if return_iters else None
    nb = lib().fiode_certify_workspace_bytes(M, n_chunks)
    ws = _ws(nb, s.device)
    check(lib().fiode_certify_grid(C.byref(dw), ptr(gridz), Ng, T, ptr(lut), ptr(s), ptr(labels), M, batches, float(kappa),
                                   ptr(out), ptr(iters), ptr(ws), nb, stream_ptr()), 'fiode_certify_grid')
    return (out, iters) if return_iters else out


class SolveResult:
    """Device-resident result of one dopri5 solve.  ``stats()`` is the only host sync."""
    ERRORS = {1: 'underflow in dt {}', 2: 'non-finite values in state `y`', 3: 'max_num_steps exceeded'}

    def __init__(self, out, stats_buf, log=None):
        self.out = out
        self._buf = stats_buf
        self._stats = None
        self._log = log            # (dts fp64, ratios fp32, capacity) device buffers of fiode_dopri5_simplex_logged

    def attempt_log(self):
        """[(dt, ratio)] of every step attempt, or None if the caller-owned log was too small (re-run larger)."""
        st = self.stats()
        n = st.n_accept + st.n_reject
        if self._log is None:
            return [(float(st.dts[i]), float(st.ratios[i])) for i in range(n)] if n <= st.n_logged else None
        dts, ratios, cap = self._log
        if n > cap:
            return None
        return list(zip(dts[:n].cpu().tolist(), ratios[:n].cpu().tolist()))

    def stats(self) -> SolveStats:
        if self._stats is None:
            raw = self._buf.cpu().numpy().tobytes()
            self._stats = SolveStats.from_buffer_copy(raw[:C.sizeof(SolveStats)])
        return self._stats

    def raise_for_status(self):
        if self._stats is None and int(self._buf.view(torch.int32)[3]) == 0:     # SolveStats.status: a 4-byte read
            return self
        st = self.stats()
        if st.status:   # same exception type and message as torchdiffeq 0.2.2
            raise AssertionError(self.ERRORS[st.status].format(st.last_dt))
        return self


def dopri5_simplex(dw: DynWeights, h0, s, t, rtol, atol, max_num_steps=2 ** 31 - 1, rows_per_cta=0, log_capacity=0):
    """Device-resident adaptive dopri5 over the simplex dynamics (replaces torchdiffeq.odeint at
    models.py:236-241).  h0 (B,10); s (1|B,10); t 1-D fp64 (device).  Returns SolveResult with
    out (len(t), B, 10).  ``log_capacity`` > 0 keeps a device log of that many step attempts
    (``SolveResult.attempt_log``); the stats struct alone holds the first 64."""
    h0, s = _f32(h0), _f32(s)
    B = h0.shape[0]
    t = t.detach().to(device=h0.device, dtype=torch.float64).contiguous()
    nt = t.numel()
    out = torch.empty(nt, B, NH, dtype=torch.float32, device=h0.device)
    stats = torch.zeros(C.sizeof(SolveStats), dtype=torch.uint8, device=h0.device)
    nb = lib().fiode_dopri5_workspace_bytes(B)
    ws = _ws(nb, h0.device)
    log = None
    if log_capacity > 0:
        log = (torch.zeros(log_capacity, dtype=torch.float64, device=h0.device),
               torch.zeros(log_capacity, dtype=torch.float32, device=h0.device), int(log_capacity))
    check(lib().fiode_dopri5_simplex_logged(C.byref(dw), ptr(h0), ptr(s), s.shape[0], B, ptr(t), nt, float(rtol), float(atol),
                                            int(max_num_steps), ptr(out), ptr(stats), ptr(ws), nb, int(rows_per_cta),
                                            ptr(log[0]) if log else None, ptr(log[1]) if log else None,
                                            int(log_capacity), stream_ptr()), 'fiode_dopri5_simplex_logged')
    return SolveResult(out, stats, log)


# ---- forward-invariance training workload ---------------------------------------------------------------
SAMPLER_KINDS = dict(uniform=0, correct_cone=1, decision_boundary=2)


def sampler_fill(kind, e, y, h, s0):
    """Write n samples per image into h (B,S,10) at sample offset s0 from Exp(1) variates e."""
    e = _f32(e)
    B, S = h.shape[0], h.shape[1]
    n = e.shape[0] if kind == 'uniform' else e.shape[1]
    yy = y.to(torch.int32).contiguous() if y is not None else None
    check(lib().fiode_sampler_fill(SAMPLER_KINDS[kind], ptr(e), ptr(yy), B, n, S, s0, ptr(h), stream_ptr()),
          'fiode_sampler_fill')


def decision_boundary(h, y, on_simplex=True, log_mode=False):
    """lya_cands.py:72-94: V = 1 + max_{j != y} p_j - p_y per row (its log in log_mode)."""
    h = _f32(h)
    yy = y.to(torch.int32).contiguous()
    V = torch.empty(h.shape[0], dtype=torch.float32, device=h.device)
    check(lib().fiode_decision_boundary(ptr(h), ptr(yy), h.shape[0], int(bool(on_simplex)), int(bool(log_mode)), ptr(V), stream_ptr()),
          'fiode_decision_boundary')
    return V


class FiSaved:
    """Device tensors the FI-loss forward keeps for its backward."""
    __slots__ = ('h', 's', 'y', 'B', 'S', 'kappa', 'm1', 'm2', 'f', 'ft', 'mu', 'z1', 'z2', 'V', 'Vdot', 'out4', 'dw', 'opt')

    def __init__(self):
        self.opt = None


class FiOpt:
    """fiode_fi_options: kappa as a device scalar (schedules under CUDA-graph replay) and in-kernel Philox dropout."""

    def __init__(self, kappa_dev=None, dropout_p=0.0, seed=0, step_dev=None):
        from ._lib import FiOptions
        self.kappa_dev, self.step_dev = kappa_dev, step_dev            # keep the tensors alive
        assert kappa_dev is None or (kappa_dev.dtype == torch.float32 and kappa_dev.numel() == 1)
        assert step_dev is None or (step_dev.dtype == torch.int64 and step_dev.numel() == 1)
        assert dropout_p == 0.0 or step_dev is not None, 'in-kernel dropout needs a device step counter'
        self.c = FiOptions(ptr(kappa_dev), float(dropout_p), int(seed) & (2 ** 64 - 1), ptr(step_dev))

    @property
    def dropout_p(self):
        return self.c.dropout_p


def dropout_mask(opt: 'FiOpt', layer, N):
    """The (N,128) mask the next forward call draws for ``layer`` (0 | 1) with these options (host sync)."""
    out = torch.empty(N, 128, dtype=torch.float32, device=opt.step_dev.device)
    check(lib().fiode_dropout_mask(opt.c.dropout_p, opt.c.seed, ptr(opt.step_dev), int(layer), N, ptr(out), stream_ptr()),
          'fiode_dropout_mask')
    return out


def fi_loss_fwd(dw, h, s, y, S, kappa, m1=None, m2=None, opt: 'FiOpt' = None):
    """loss = mean relu(Vdot + kappa V) over the B*S samples (pl_modules.py:435-466).  Returns FiSaved;
    ``out4`` = [loss, effective_batch_size, mean_active_constraints, bisection iterations] on the device."""
    h, s = _f32(h), _f32(s)
    B = s.shape[0]
    N = B * S
    assert h.shape == (N, NH)
    dev = h.device
    sv = FiSaved()
    sv.h, sv.s, sv.y, sv.B, sv.S, sv.kappa, sv.dw = h, s, y.to(torch.int32).contiguous(), B, S, float(kappa), dw
    sv.m1 = _f32(m1) if m1 is not None else None
    sv.m2 = _f32(m2) if m2 is not None else None
    new = lambda *shape: torch.empty(*shape, dtype=torch.float32, device=dev)
    sv.f, sv.ft, sv.mu, sv.z1, sv.z2, sv.V, sv.Vdot = new(N, NH), new(N, NH), new(N), new(N, 128), new(N, 128), new(N), new(N)
    sv.out4 = new(4)
    nb = lib().fiode_fi_loss_workspace_bytes(N, B)
    ws = _ws(nb, dev)
    sv.opt = opt
    check(lib().fiode_fi_loss_fwd_opt(C.byref(dw), ptr(h), ptr(s), ptr(sv.y), B, S, sv.kappa, ptr(sv.m1), ptr(sv.m2),
                                      C.byref(opt.c) if opt else None, ptr(sv.f),
                                      ptr(sv.ft), ptr(sv.mu), ptr(sv.z1), ptr(sv.z2), ptr(sv.V), ptr(sv.Vdot), ptr(sv.out4),
                                      ptr(ws), nb, stream_ptr()), 'fiode_fi_loss_fwd_opt')
    return sv


def fi_loss_bwd(sv: FiSaved, grad_loss=1.0):
    """Gradients of the FI loss w.r.t. Q1, Qx, Q2, Q3, b1(=bx), b2, b3 and s."""
    dev = sv.h.device
    new = lambda *shape: torch.empty(*shape, dtype=torch.float32, device=dev)
    g = dict(Q1=new(128, NH), Qx=new(128, NH), Q2=new(128, 128), Q3=new(NH, 128), b1=new(128), b2=new(128), b3=new(NH),
             s=new(sv.B, NH))
    N = sv.B * sv.S
    nb = lib().fiode_fi_loss_workspace_bytes(N, sv.B)
    ws = _ws(nb, dev)
    opt = getattr(sv, 'opt', None)
    check(lib().fiode_fi_loss_bwd_opt(C.byref(sv.dw), ptr(sv.h), ptr(sv.s), ptr(sv.y), sv.B, sv.S, sv.kappa, float(grad_loss),
                                      ptr(sv.m1), ptr(sv.m2), C.byref(opt.c) if opt else None, ptr(sv.f), ptr(sv.ft), ptr(sv.mu),
                                      ptr(sv.z1), ptr(sv.z2),
                                      ptr(sv.V), ptr(sv.Vdot), ptr(g['Q1']), ptr(g['Qx']), ptr(g['Q2']), ptr(g['Q3']),
                                      ptr(g['b1']), ptr(g['b2']), ptr(g['b3']), ptr(g['s']), ptr(ws), nb, stream_ptr()),
          'fiode_fi_loss_bwd_opt')
    return g


class BatchedSolveResult:
    """Result of G independent solves in one launch.  ``stats()`` -> list of SolveStats (one host sync)."""

    def __init__(self, out, stats_buf, G):
        self.out, self._buf, self.G, self._stats = out, stats_buf, G, None

    def stats(self):
        if self._stats is None:
            raw = self._buf.cpu().numpy().tobytes()
            n = C.sizeof(SolveStats)
            self._stats = [SolveStats.from_buffer_copy(raw[g * n:(g + 1) * n]) for g in range(self.G)]
        return self._stats

    def status_max(self):
        """Largest status word of the G solves, on the device (0 = every solve finished)."""
        return self._buf.view(torch.int32).view(self.G, -1)[:, 3].max()             # SolveStats.status

    def raise_for_status(self):
        """The end-of-solve check (torchdiffeq's assertions): ONE 4-byte read -- the largest status word of the G solves --
        and the full statistics only when it is non-zero."""
        if self._stats is None:
            if int(self.status_max()) == 0:
                return self
        for st in self.stats():
            if st.status:
                raise AssertionError(SolveResult.ERRORS[st.status].format(st.last_dt))
        return self


def dopri5_simplex_batched(dw: DynWeights, h0, s, G, t, rtol, atol, max_num_steps=2 ** 31 - 1, cluster_size=0):
    """G independent batch solves (rows g*B .. (g+1)*B-1 form solve g) in ONE launch, a thread-block
    cluster per solve: each keeps its own batch-wide error norm and projection stop rule, so solve g is
    what ``dopri5_simplex`` returns for its rows alone.  h0 (G*B,10); s (1|G*B,10).  out (len(t), G*B, 10)."""
    h0, s = _f32(h0), _f32(s)
    N = h0.shape[0]
    assert N % G == 0
    B = N // G
    t = t.detach().to(device=h0.device, dtype=torch.float64).contiguous()
    nt = t.numel()
    out = torch.empty(nt, N, NH, dtype=torch.float32, device=h0.device)
    stats = torch.zeros(G * C.sizeof(SolveStats), dtype=torch.uint8, device=h0.device)
    check(lib().fiode_dopri5_simplex_batched(C.byref(dw), ptr(h0), ptr(s), s.shape[0], G, B, ptr(t), nt, float(rtol),
                                             float(atol), int(max_num_steps), ptr(out), ptr(stats), int(cluster_size),
                                             stream_ptr()), 'fiode_dopri5_simplex_batched')
    return BatchedSolveResult(out, stats, G)


def dopri5_max_concurrent_solves(B, cluster_size=0):
    """Solves of B rows one ``dopri5_simplex_batched`` launch keeps co-resident (a cluster each); more run in waves."""
    n = lib().fiode_dopri5_max_concurrent_solves(int(B), int(cluster_size))
    if n < 0:
        check(n, 'fiode_dopri5_max_concurrent_solves')
    return n


# ---- training: adjoints of the backbone layers and of the Cayley transform ------------------------
def _opt(t):
    return ptr(t) if t is not None else None


def normalize(x, mu, std):
    """Normalize.forward (models.py:23-26)."""
    x = _f32(x)
    B, Cc = x.shape[:2]
    inner = x[0, 0].numel()
    y = torch.empty_like(x)
    check(lib().fiode_normalize(ptr(x), B, Cc, inner, ptr(_f32(mu).reshape(-1)), ptr(_f32(std).reshape(-1)), ptr(y),
                                stream_ptr()), 'fiode_normalize')
    return y


def group_sort_bwd(x_pre, gy):
    x_pre, gy = _f32(x_pre), _f32(gy)
    B, Cc = x_pre.shape[:2]
    inner = x_pre[0, 0].numel() if x_pre.dim() > 2 else 1
    gx = torch.empty_like(x_pre)
    check(lib().fiode_group_sort_bwd(ptr(x_pre), ptr(gy), B, Cc, inner, ptr(gx), stream_ptr()), 'fiode_group_sort_bwd')
    return gx


def linear_bwd(X, Q, gY, need_gx=True, need_gb=True):
    """Adjoint of linear_fwd: returns (gX or None, gQ, gb or None)."""
    X, Q, gY = _f32(X), _f32(Q), _f32(gY)
    X2, g2 = X.reshape(-1, X.shape[-1]), gY.reshape(-1, gY.shape[-1])
    M, K = X2.shape
    N = Q.shape[0]
    gX = torch.empty(M, K, dtype=torch.float32, device=X.device) if need_gx else None
    gQ = torch.empty(N, K, dtype=torch.float32, device=X.device)
    gb = torch.empty(N, dtype=torch.float32, device=X.device) if need_gb else None
    nb = lib().fiode_linear_bwd_workspace_bytes(M, N, K)
    ws = _ws(nb, X.device)
    check(lib().fiode_linear_bwd(ptr(X2), ptr(Q), ptr(g2), M, N, K, _opt(gX), ptr(gQ), _opt(gb), ptr(ws), nb, stream_ptr()),
          'fiode_linear_bwd')
    return (gX.reshape(X.shape) if need_gx else None), gQ, gb


def conv1x1_s2d_bwd(x, Q, gy, need_gx=True, need_gb=True):
    x, Q, gy = _f32(x), _f32(Q), _f32(gy)
    B, cin, n, _ = x.shape
    cout = Q.shape[0]
    gx = torch.empty_like(x) if need_gx else None
    gQ = torch.empty_like(Q)
    gb = torch.empty(cout, dtype=torch.float32, device=x.device) if need_gb else None
    nb = lib().fiode_conv1x1_s2d_bwd_workspace_bytes(B, cin, cout, n)
    ws = _ws(nb, x.device)
    check(lib().fiode_conv1x1_s2d_bwd(ptr(x), ptr(Q), ptr(gy), B, cin, cout, n, _opt(gx), ptr(gQ), _opt(gb), ptr(ws), nb,
                                      stream_ptr()), 'fiode_conv1x1_s2d_bwd')
    return gx, gQ, gb


def cayley_conv_layer_bwd(x, Qf, gy, need_gx=True, need_gb=True):
    """Adjoint of cayley_conv_fwd: returns (gx or None, gQf complex (F,cout,cin), gb or None)."""
    x, gy = _f32(x), _f32(gy)
    B, cin, n, _ = x.shape
    Qr = torch.view_as_real(Qf).contiguous()
    F, cout = Qr.shape[:2]
    gx = torch.empty_like(x) if need_gx else None
    gQ = torch.empty_like(Qr)
    gb = torch.empty(cout, dtype=torch.float32, device=x.device) if need_gb else None
    nb = lib().fiode_cayley_conv_layer_bwd_workspace_bytes(B, cin, cout, n)
    ws = _ws(nb, x.device)
    check(lib().fiode_cayley_conv_layer_bwd(ptr(x), ptr(Qr), ptr(gy), B, cin, cout, n, _opt(gx), ptr(gQ), _opt(gb), ptr(ws),
                                            nb, stream_ptr()), 'fiode_cayley_conv_layer_bwd')
    return gx, torch.view_as_complex(gQ), gb


def cayley_linear_bwd(weight, alpha, gQ, minv=None, ctx=None):
    """Adjoint of cayley_linear: (gW, galpha).  ``minv``: (I + A)^-1 from ``cayley_linear(..., save_minv=True)``; ``ctx``: the
    context buffer from ``cayley_linear(..., save_ctx=True)``."""
    W, gQ = _f32(weight), _f32(gQ)
    cout, cin = W.shape
    a = _f32(alpha).reshape(1)
    gW = torch.empty_like(W)
    ga = torch.empty(1, dtype=torch.float32, device=W.device)
    nb = lib().fiode_cayley_linear_bwd_workspace_bytes(cout, cin)
    ws = _ws(nb, W.device)
    if ctx is not None:
        check(lib().fiode_cayley_linear_bwd_ctx(ptr(W), ptr(a), cout, cin, ptr(gQ), ptr(ctx), ptr(gW), ptr(ga), ptr(ws), nb,
                                                stream_ptr()), 'fiode_cayley_linear_bwd_ctx')
        return gW, ga
    if minv is not None:
        check(lib().fiode_cayley_linear_bwd_saved(ptr(W), ptr(a), cout, cin, ptr(gQ), ptr(_f32(minv)), ptr(gW), ptr(ga), ptr(ws), nb,
                                                  stream_ptr()), 'fiode_cayley_linear_bwd_saved')
        return gW, ga
    check(lib().fiode_cayley_linear_bwd(ptr(W), ptr(a), cout, cin, ptr(gQ), ptr(gW), ptr(ga), ptr(ws), nb, stream_ptr()),
          'fiode_cayley_linear_bwd')
    return gW, ga


def cayley_small_ok(cout, cin):
    """True when the layer's Cayley transform runs as one CTA (and so can ride in a cayley_small_batch launch)."""
    return bool(lib().fiode_cayley_small_ok(int(cout), int(cin)))


def _ptr_list(ts):
    arr = (C.c_void_p * len(ts))()
    for i, t in enumerate(ts):
        arr[i] = None if t is None else ptr(t).value
    return arr


def _small_batch_args(weights, alphas, mults):
    Ws = [_f32(w).reshape(w.shape[0], w.shape[1]) for w in weights]          # (cout, cin[, 1, 1]) -> (cout, cin)
    As = [_f32(a).reshape(1) for a in alphas]
    k = len(Ws)
    couts = (C.c_int * k)(*[w.shape[0] for w in Ws])
    cins = (C.c_int * k)(*[w.shape[1] for w in Ws])
    mu = (C.c_float * k)(*[float(m) for m in mults])
    return Ws, As, k, couts, cins, mu


def cayley_small_batch(weights, alphas, mults):
    """Q_i = cayley(alpha_i W_i / (||W_i|| mult_i)) for several small layers in ONE launch (one CTA per layer).  Returns
    ([Q_i], [Minv_i]); bit-identical with ``cayley_linear`` / ``cayley_conv`` (k = 1, mult = sqrt(F)) of each layer."""
    Ws, As, k, couts, cins, mu = _small_batch_args(weights, alphas, mults)
    Qs = [torch.empty_like(w) for w in Ws]
    Ms = [torch.empty(min(w.shape), min(w.shape), dtype=torch.float32, device=w.device) for w in Ws]
    check(lib().fiode_cayley_small_batch_fwd(k, _ptr_list(Ws), _ptr_list(As), couts, cins, mu, _ptr_list(Qs), _ptr_list(Ms),
                                             stream_ptr()), 'fiode_cayley_small_batch_fwd')
    return Qs, Ms


def cayley_small_batch_bwd(weights, alphas, mults, gQs, minvs):
    """Adjoint of ``cayley_small_batch``: ([gW_i] shaped like W_i, [galpha_i] (1,)), one launch."""
    Ws, As, k, couts, cins, mu = _small_batch_args(weights, alphas, mults)
    Gs = [_f32(g).reshape(w.shape) for g, w in zip(gQs, Ws)]
    Ms = [None if m is None else _f32(m) for m in minvs]
    gWs = [torch.empty_like(w) for w in Ws]
    gas = [torch.empty(1, dtype=torch.float32, device=w.device) for w in Ws]
    check(lib().fiode_cayley_small_batch_bwd(k, _ptr_list(Ws), _ptr_list(As), couts, cins, mu, _ptr_list(Gs), _ptr_list(Ms),
                                             _ptr_list(gWs), _ptr_list(gas), stream_ptr()), 'fiode_cayley_small_batch_bwd')
    return [g.reshape(w.shape) for g, w in zip(gWs, weights)], gas


def cayley_conv_bwd(weight, alpha, n, gQf):
    """Adjoint of cayley_conv: (gw (cout,cin,k,k), galpha)."""
    w = _f32(weight)
    cout, cin, k, _ = w.shape
    a = _f32(alpha).reshape(1)
    g = (torch.view_as_real(gQf) if gQf.is_complex() else gQf).to(torch.float32).contiguous()
    gw = torch.empty_like(w)
    ga = torch.empty(1, dtype=torch.float32, device=w.device)
    nb = lib().fiode_cayley_conv_bwd_workspace_bytes(cout, cin, k, n)
    ws = _ws(nb, w.device)
    check(lib().fiode_cayley_conv_bwd(ptr(w), ptr(a), cout, cin, k, n, ptr(g), ptr(gw), ptr(ga), ptr(ws), nb, stream_ptr()),
          'fiode_cayley_conv_bwd')
    return gw, ga


# ---- interval step after the CROWN bound (certify_crown.py:139-150) ----------------------------------
def ibp_sigmoid(lb, ub, h, eps, alpha_1, alpha_2, sigma_1):
    lb, ub, h = _f32(lb), _f32(ub), _f32(h)
    N = h.shape[0]
    o_lb, o_ub = torch.empty_like(lb), torch.empty_like(ub)
    check(lib().fiode_ibp_sigmoid(ptr(lb), ptr(ub), ptr(h), N, float(eps), float(alpha_1), float(alpha_2), float(sigma_1),
                                  ptr(o_lb), ptr(o_ub), stream_ptr()), 'fiode_ibp_sigmoid')
    return o_lb, o_ub


def ibp_cbf_qp(h, eps, lb, ub, alpha_1, sigma_1, max_iter=30, tol=1e-4, return_iters=False):
    h, lb, ub = _f32(h), _f32(lb), _f32(ub)
    N = h.shape[0]
    f_lb, f_ub = torch.empty_like(h), torch.empty_like(h)
    it = torch.zeros(2, dtype=torch.int32, device=h.device)
    ws = _ws(lib().fiode_ibp_workspace_bytes(), h.device)
    check(lib().fiode_ibp_cbf_qp(ptr(h), ptr(lb), ptr(ub), N, float(eps), float(alpha_1), float(sigma_1), int(max_iter),
                                 float(tol), ptr(f_lb), ptr(f_ub), ptr(it), ptr(ws), ws.numel(), stream_ptr()), 'fiode_ibp_cbf_qp')
    return (f_lb, f_ub, it) if return_iters else (f_lb, f_ub)


def perturbed_vdot(f_lb, f_ub, label, runner_up, return_max=False):
    f_lb, f_ub = _f32(f_lb), _f32(f_ub)
    N = f_lb.shape[0]
    lab = label.detach().to(device=f_lb.device, dtype=torch.int32).reshape(-1).contiguous()
    ru = runner_up.detach().to(device=f_lb.device, dtype=torch.uint8).contiguous()
    vdot = torch.empty(N, dtype=torch.float32, device=f_lb.device)
    vmax = torch.empty(1, dtype=torch.float32, device=f_lb.device) if return_max else None
    ws = _ws(256, f_lb.device)
    check(lib().fiode_perturbed_vdot(ptr(f_lb), ptr(f_ub), ptr(lab), lab.numel(), ptr(ru), N, ptr(vdot), _opt(vmax), ptr(ws),
                                     ws.numel(), stream_ptr()), 'fiode_perturbed_vdot')
    return (vdot, vmax) if return_max else vdot


def runner_up_mask(eta, label, margin):
    eta = _f32(eta)
    N = eta.shape[0]
    mask = torch.empty(N, NH, dtype=torch.uint8, device=eta.device)
    check(lib().fiode_runner_up_mask(ptr(eta), N, int(label), float(margin), ptr(mask), stream_ptr()), 'fiode_runner_up_mask')
    return mask


# ---- differentiable solve: right-hand-side VJPs (csrc/ode_bwd.cu) ---------------------------------------
class RhsEval:
    """One evaluation of f(h, s) with what its adjoint needs."""
    __slots__ = ('h', 'f', 'ft', 'mu', 'z1', 'z2', 'dft', 'd2', 'd1')


def rhs_fwd_train(dw: DynWeights, h, s):
    h, s = _f32(h), _f32(s)
    N = h.shape[0]
    assert s.shape[0] == N, 'one static-feature row per state row'
    ev = RhsEval()
    ev.h = h
    ev.f, ev.ft = torch.empty_like(h), torch.empty_like(h)
    ev.mu = torch.empty(N, dtype=torch.float32, device=h.device)
    ev.z1 = torch.empty(N, 128, dtype=torch.float32, device=h.device)
    ev.z2 = torch.empty(N, 128, dtype=torch.float32, device=h.device)
    nb = lib().fiode_rhs_train_workspace_bytes(N)
    ws = _ws(nb, h.device)
    check(lib().fiode_rhs_fwd_train(C.byref(dw), ptr(h), ptr(s), N, ptr(ev.f), ptr(ev.ft), ptr(ev.mu), ptr(ev.z1), ptr(ev.z2),
                                    ptr(ws), nb, stream_ptr()), 'fiode_rhs_fwd_train')
    return ev


def dyn_transposes(dw: DynWeights, device):
    t = [torch.empty(10, 128, device=device), torch.empty(10, 128, device=device), torch.empty(128, 128, device=device),
         torch.empty(128, 10, device=device)]
    check(lib().fiode_dyn_transposes(C.byref(dw), *[ptr(x) for x in t], stream_ptr()), 'fiode_dyn_transposes')
    return t


def rhs_bwd_data(dw: DynWeights, ev: RhsEval, g_f, QT):
    """g_f -> (g_h, g_s); the layer deltas are stored on ``ev`` for the stacked weight-gradient GEMMs."""
    g_f = _f32(g_f)
    N = ev.h.shape[0]
    dev = ev.h.device
    ev.dft = torch.empty(N, 10, dtype=torch.float32, device=dev)
    ev.d2 = torch.empty(N, 128, dtype=torch.float32, device=dev)
    ev.d1 = torch.empty(N, 128, dtype=torch.float32, device=dev)
    g_h, g_s = torch.empty(N, 10, dtype=torch.float32, device=dev), torch.empty(N, 10, dtype=torch.float32, device=dev)
    nb = lib().fiode_rhs_train_workspace_bytes(N)
    ws = _ws(nb, dev)
    check(lib().fiode_rhs_bwd_data(C.byref(dw), ptr(ev.h), N, ptr(g_f), ptr(ev.f), ptr(ev.ft), ptr(ev.mu), ptr(ev.z1), ptr(ev.z2),
                                   ptr(QT[0]), ptr(QT[1]), ptr(QT[2]), ptr(QT[3]), ptr(ev.dft), ptr(ev.d2), ptr(ev.d1),
                                   ptr(g_h), ptr(g_s), ptr(ws), nb, stream_ptr()), 'fiode_rhs_bwd_data')
    return g_h, g_s


def rhs_bwd_params(h, s, z1, z2, dft, d2, d1):
    M = h.shape[0]
    dev = h.device
    o = dict(Q1=torch.empty(128, 10, device=dev), Qx=torch.empty(128, 10, device=dev), Q2=torch.empty(128, 128, device=dev),
             Q3=torch.empty(10, 128, device=dev), b1=torch.empty(128, device=dev), b2=torch.empty(128, device=dev),
             b3=torch.empty(10, device=dev))
    nb = lib().fiode_rhs_bwd_params_workspace_bytes(M)
    ws = _ws(nb, dev)
    check(lib().fiode_rhs_bwd_params(ptr(h), ptr(s), ptr(z1), ptr(z2), ptr(dft), ptr(d2), ptr(d1), M, ptr(o['Q1']), ptr(o['Qx']),
                                     ptr(o['Q2']), ptr(o['Q3']), ptr(o['b1']), ptr(o['b2']), ptr(o['b3']), ptr(ws), nb,
                                     stream_ptr()), 'fiode_rhs_bwd_params')
    return o


def lincomb(out, terms):
    """out = sum c_i x_i for ``terms`` = [(c, x), ...] (at most 8; out may be one of the x)."""
    n = len(terms)
    assert 1 <= n <= 8
    xs = (C.c_void_p * n)(*[t[1].data_ptr() for t in terms])
    cs = (C.c_float * n)(*[float(t[0]) for t in terms])
    for _, x in terms:
        assert x.is_cuda and x.is_contiguous() and x.dtype == torch.float32 and x.numel() == out.numel()
    check(lib().fiode_lincomb(ptr(out), out.numel(), n, xs, cs, stream_ptr()), 'fiode_lincomb')
    return out


# ---- every op runs on the device of its tensors --------------------------------------------------------------------
# The C ABI launches on the CURRENT CUDA device (per-device function attributes, SM counts, the current stream); a
# process that drives several GPUs must therefore make the tensors' device current around each call.
def _first_cuda_device(args, kwargs):
    for a in list(args) + list(kwargs.values()):
        if isinstance(a, torch.Tensor) and a.is_cuda:
            return a.device
        for name in ('h', 'out', 'step_dev', 'kappa_dev', '_keep'):
            t = getattr(a, name, None)
            if isinstance(t, torch.Tensor) and t.is_cuda:
                return t.device
            if isinstance(t, dict):
                for v in t.values():
                    if isinstance(v, torch.Tensor) and v.is_cuda:
                        return v.device
    return None


def _on_tensor_device(fn):
    import functools

    @functools.wraps(fn)
    def wrapped(*args, **kwargs):
        dev = _first_cuda_device(args, kwargs)
        if dev is None or dev.index == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(dev):
            return fn(*args, **kwargs)
    return wrapped


for _name, _obj in list(globals().items()):
    if callable(_obj) and getattr(_obj, '__module__', None) == __name__ and not _name.startswith('_') \
            and not isinstance(_obj, type):
        globals()[_name] = _on_tensor_device(_obj)
