#!/usr/bin/env python
"""bench.py -- headline benchmark of the FI-ODE hot path on B200 (contract: see README / DESIGN.md).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): CIFAR-10 Cayley ortho-conv neural ODE inference, dopri5
rtol=atol=1e-3, t_max=1, batch 256, synthetic 3x32x32 inputs, random-init weights.  One *step* is one
pass of the hot path over ``--solves-per-step`` independent batch-256 solves (each keeps its own
batch-wide RMS error norm and batch-global projection stop rule, exactly like one reference call);
metric = batch-256 ODE solves per second, whole job.  A secondary ``certify`` object reports the
certify_lipschitz grid-check throughput (grid points x images per second).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = 'ODE-solves/s (CIFAR dopri5 b=256)'
UNIT = 'batch-256 solves/s'
B = 256
TOL = 1e-3
T_MAX = 1.0
MU = (0.485, 0.456, 0.406)
STD = (0.225, 0.225, 0.225)
RHS_FLOP = 37888                     # per row per RHS evaluation
BACKBONE_FLOP = 1.31e7               # per image, minimal algorithm
FFMA_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12     # fp32 FFMA pipe at max clock, for context


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d['hbm_gbs'], bf16=d['bf16_tflops'], bf16_sustained=d.get('bf16_tflops_sustained'),
                    source='measured')
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, source='fallback')


# ---- synthetic model ------------------------------------------------------------------------------
def build_model(device, scale_nominal=False):
    """Random-init weights of the named architecture under torch.manual_seed(0), built through the
    fiode_b200 module tree (same constructors / init distributions as the reference classes)."""
    from fiode_b200 import models
    from fiode_b200.dynamics.classification import OrthoClassDynProjectSimplexLips
    from fiode_b200.dynamics.init_coordinates import UniformInitFun
    torch.manual_seed(0)
    param_map = models.make_ortho_KWLarge_Concat(3, 10, mu=MU, std=STD, out_dim=10, act='GroupSort')
    dyn = OrthoClassDynProjectSimplexLips(n_hidden=10, activation='ReLU', mlp_size=128, alpha_1=100., alpha_2=20.,
                                          sigma_1=0.02, x_dim=10, cayley=True, dropout=0.5,
                                          scale_nominal=scale_nominal, kappa=2.0, kappa_length=0)
    ivp = models.IVP(3072, 10, dyn, UniformInitFun((10,), param_map=param_map), ts=torch.linspace(0., T_MAX, 2))
    return ivp.to(device).eval()


def cpu_state_dict(ivp):
    """Reference-layout state_dict (CPU) of the benchmark model, for the oracle / CPU baseline."""
    return {'model.' + k: v.detach().cpu() for k, v in ivp.state_dict().items() if v is not None}


class ClockSampler:
    QUERY = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
             'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index = index
        self.f = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(['nvidia-smi', f'--query-gpu={self.QUERY}', '--format=csv,noheader,nounits',
                                       '-i', str(self.index), '-lms', '100'], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['nvidia-smi unavailable'])
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [r.split(',') for r in open(self.f.name).read().strip().splitlines() if r.count(',') >= 8]
        os.unlink(self.f.name)
        if not rows:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['no samples'])
        sm = sorted(float(r[1]) for r in rows)
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for i, n in enumerate(names) if any('Active' in r[5 + i] and 'Not' not in r[5 + i] for r in rows)]
        return dict(sm_mhz=sm[len(sm) // 2], sm_max_mhz=float(rows[0][2]), reasons=reasons, samples=len(rows),
                    power_w_max=max(float(r[3]) for r in rows))


# ---- reference arm (CPU) ----------------------------------------------------------------------------
def cpu_solve_rate(sd, n_solves, threads):
    """The oracle port of the reference path (backbone + restated torchdiffeq dopri5) on host cores."""
    from oracle import dynamics as od
    torch.set_num_threads(threads)
    cfg = dict(alpha_1=100., alpha_2=20., sigma_1=0.02, scale_nominal=False)
    g = torch.Generator().manual_seed(1234)
    x = torch.rand(B, 3, 32, 32, generator=g)
    od.ivp_forward(x, sd, cfg, t_max=T_MAX, tol=TOL)          # warm-up
    t0 = time.perf_counter()
    for _ in range(n_solves):
        od.ivp_forward(x, sd, cfg, t_max=T_MAX, tol=TOL)
    dt = time.perf_counter() - t0
    return n_solves / dt, dt


def reference_state_dict():
    """Weights for the CPU arm without touching the GPU: same constructors, same seed."""
    from tests import fixtures as fx
    return fx.synth_state_dict()


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sd = reference_state_dict()
    per_step = 2
    t_all = []
    from oracle import dynamics as od
    torch.set_num_threads(threads)
    cfg = dict(alpha_1=100., alpha_2=20., sigma_1=0.02, scale_nominal=False)
    x = torch.rand(B, 3, 32, 32, generator=torch.Generator().manual_seed(1234))
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        for _ in range(per_step):
            od.ivp_forward(x, sd, cfg, t_max=T_MAX, tol=TOL)
        if i >= args.warmup:
            t_all.append(time.perf_counter() - t0)
    total = sum(t_all)
    value = per_step * args.steps / total
    line = dict(metric=METRIC, value=value, unit=UNIT, impl='reference', n_gpus=args.gpus, steps=args.steps,
                warmup=args.warmup, ms_per_step=1e3 * total / args.steps, higher_is_better=True, scaling='weak',
                vs_baseline=None, dtype='f32', data='synthetic',
                config=dict(workload='cifar_cayley_node_inference_dopri5_b256_tol1e-3_tmax1', batch=B,
                            solves_per_step=per_step, note='oracle port of the reference path on host cores '
                            '(reference checkout and torchdiffeq absent on the GPU box)'),
                cpu_baseline=dict(value=value, unit=UNIT, cores=threads, kind='port',
                                  sample=f'{per_step} batch-256 solves per step x {args.steps} steps'),
                e2e=dict(value=value, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line))


# ---- B200 arm -------------------------------------------------------------------------------------------
def timed(fn, iters, stream=None):
    """CUDA-event time (ms) of ``iters`` calls of fn on the current stream."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


def certify_bench(ivp, device, pk, world=1, iters=2):
    """Secondary metric (BASELINE config 4): certify_lipschitz grid check over grid_40 (T=40: 41 320 837 decision-boundary
    points per label), M images per rank (images shard across ranks with no collective; weak scaling), L2 norm, eps 0.141."""
    from fiode_b200 import ops, grid as fgrid
    import fiode_b200
    T, M = 40, 148
    rank = int(os.environ.get('RANK', '0'))
    gz = fgrid.boundary_grid_numerators(10, T, device=device)
    g = torch.Generator().manual_seed(1236 + rank)
    s = (torch.randn(M, 10, generator=g) * 0.5).to(device)
    labels = torch.randint(0, 10, (M,), generator=g).to(device)
    kappa = math.sqrt(2) * (1 / 0.225) * 0.141
    dw = ivp.dyn_fun.dyn_weights()
    fn = lambda: ops.certify_grid(dw, gz, T, s, labels, kappa, batches=10)
    fn()
    c0 = fiode_b200.launch_count()
    ms = timed(fn, iters) / iters
    launches = (fiode_b200.launch_count() - c0) // iters
    t = torch.tensor([ms], device=device)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    units = world * M * gz.shape[0]
    tf = units * RHS_FLOP / (ms * 1e-3) / 1e12 / world          # per GPU, for the roofline
    return dict(metric='certified grid pts/s (certify_lipschitz, image x grid-point pairs)', value=units / (ms * 1e-3),
                unit='grid pts/s (whole job)', workload=f'grid_{T} ({gz.shape[0]} pts/label) x {M} images per GPU x {world} GPU, L2 norm, batches=10',
                ms=ms, gpu_launches=launches, full_config_seconds=10000 * gz.shape[0] / (units / (ms * 1e-3)),
                roofline=dict(bound='tensor', achieved=tf, peak=pk['bf16'], unit='TFLOP/s', frac=tf / pk['bf16'],
                              traffic=None, peak_source=pk['source'],
                              kernel='certify_tc_kernel',
                              note='per GPU; tcgen05, fp16x2-split FP32 emulation (3 MMAs per product: ceiling = peak / 3, i.e. frac %.3f '
                                   'of that ceiling); %.2fx the 74.4 TFLOP/s FP32 FFMA pipe the reference formulation would '
                                   'be bound by; ncu: issue-bound (profiles/)' % (3 * tf / pk['bf16'], tf / FFMA_PEAK_TFLOPS)))


def fi_step_bench(device, pk, world, iters=10, warm=8):
    """Secondary metric (BASELINE config 3): one data-parallel FI training step = sampler + backbone forward (train
    mode) + fused FI loss (B=128, S=h_sample_size=256 -> 32 768 rows, dropout 0.5) + loss.backward() through every
    adjoint down to the 33 parameters + ONE flat-buffer gradient all-reduce (NCCL when world > 1).  No optimizer
    step: Adam stays with the reference's Lightning loop (pl_modules.py:506-520)."""
    import fiode_b200
    from fiode_b200 import models, pl_modules, lya_cands
    from fiode_b200.sampling import sampler as smp, sampler_schedulers as sch
    from fiode_b200.dynamics.classification import OrthoClassDynProjectSimplexLips
    from fiode_b200.dynamics.init_coordinates import UniformInitFun
    from fiode_b200.dynamics.output_coordinates import DefaultOutputFun
    from fiode_b200.distributed import FlatGradAllReduce
    Bt, S = 128, 256
    torch.manual_seed(0)
    pm = models.make_ortho_KWLarge_Concat(3, 10, mu=MU, std=STD, out_dim=10, act='GroupSort')
    dyn = OrthoClassDynProjectSimplexLips(n_hidden=10, activation='ReLU', mlp_size=128, alpha_1=100., alpha_2=20., sigma_1=0.02,
                                          x_dim=10, cayley=True, dropout=0.5, scale_nominal=True, kappa=2.0, kappa_length=0)
    mod = pl_modules.LyapunovLearning(
        order=1, h_sample_size=S, h_dist_lim=15, lya_cand=lya_cands.DecisionBoundary(on_simplex=True),
        sampler=smp.CompositeSampler((10,), [smp.UniformSimplexSampling((10,)), smp.CorrectConeSampling((10,))]),
        sampler_scheduler=sch.CompositeSamplerScheduler([sch.ConstantScheduler(0.5), sch.ConstantScheduler(0.5)], [1., 1.]),
        dynamics=dyn, output=DefaultOutputFun(), init_fun=UniformInitFun((10,), param_map=pm), t_max=T_MAX,
        train_ode_tol=TOL, val_ode_tol=TOL, epoch_off_scale=10).to(device).train()
    mod.current_epoch, mod.global_step = 0, 10 ** 9              # epoch 0 (scale_nominal on), kappa ramp finished (kappa_length=0)
    rank = int(os.environ.get('RANK', '0'))
    g = torch.Generator().manual_seed(1234 + rank)
    x = torch.rand(Bt, 3, 32, 32, generator=g).to(device)
    y = torch.randint(0, 10, (Bt,), generator=torch.Generator().manual_seed(1235 + rank)).to(device)
    mod.compute_loss(x, y, Bt, act='relu').backward()          # creates the lazily initialised conv alphas
    reducer = FlatGradAllReduce(mod.parameters())

    def compute():
        reducer.prepare()              # one memset; the backward accumulates straight into the flat all-reduce buffer
        loss = mod.compute_loss(x, y, Bt, act='relu')
        loss.backward()
        return loss

    def step():
        loss = compute()
        reducer.step()
        return loss

    for _ in range(warm):
        step()
    c0 = fiode_b200.launch_count()
    ms_eager = timed(step, iters) / iters
    launches = (fiode_b200.launch_count() - c0) // iters
    # the step is launch-bound on the host: capture it once in a CUDA graph and replay (fiode_b200/graphs.py)
    ms, graphed = ms_eager, False
    if True:
        try:
            from fiode_b200.graphs import GraphedStep
            if world == 1 or os.environ.get('FIODE_GRAPH_NCCL') == '1':
                gstep = GraphedStep(step, warmup=1)
            else:       # N > 1: forward + backward replayed from the graph, the flat NCCL all-reduce launched eagerly after it
                gcompute = GraphedStep(compute, warmup=1)

                def gstep():
                    loss = gcompute()
                    reducer.step()
                    return loss
            gstep(); gstep()
            ms, graphed = timed(gstep, iters) / iters, True
            step = gstep
        except Exception as e:           # keep the eager number
            graph_error = repr(e)
    t = torch.tensor([ms], device=device)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    rows = Bt * S
    flop = 3 * (rows * RHS_FLOP + Bt * BACKBONE_FLOP + 4.3e9)       # forward + ~2x backward
    return dict(metric='FI training steps/s (B=128 x h_sample_size=256, data-parallel)', value=world / (ms * 1e-3),
                unit='steps/s (whole job)', ms_per_step=ms, ms_per_step_eager=ms_eager, cuda_graph=graphed, gpu_launches=launches, rows_per_step=rows,
                grad_floats=reducer.numel, allreduce_bytes=4 * reducer.numel,
                algorithmic_tflops=flop / (ms * 1e-3) / 1e12, loss=float(step().detach()),
                **({} if graphed else {'graph_error': locals().get('graph_error')}),
                note='forward + backward to all 33 parameters + one flat all-reduce; kernels are the first correct versions '
                     '(FFMA GEMMs, per-layer launches), not yet tuned')


# per image: algorithmic HBM bytes (fp32 activations in + out; spectra complex64) and flops of each backbone stage
#   name, kernel, bytes/img, flop/img
STAGES = [
    ('weight packing (per call)', 'pack_w_kernel x5 + mix_real_weights', 0, 0),
    ('rfft2 32x32, 3 ch (+Normalize)', 'rfft2_kernel<32>', 3 * 1024 * 4 + 3 * 544 * 8, 0.12e6),
    ('mix 3->32 + irfft2 32x32 + bias + GroupSort', 'irfft32_mix3_wide_kernel', 3 * 544 * 8 + 32 * 1024 * 4, 0.42e6 + 1.1e6),
    ('conv 128->32 @16x16 (s2d 1x1) + GroupSort', 'gemm_tc_kernel<s2d,32>', 32 * 1024 * 4 + 32 * 256 * 4, 2 * 128 * 32 * 256),
    ('rfft2 16x16, 32 ch', 'rfft2_kernel<16,fm>', 32 * 256 * 4 + 32 * 144 * 8, 0.3e6),
    ('mixing 144 x (64->128 real)', 'gemm_tc_kernel<row,128> batched', (32 + 64) * 144 * 8, 2 * 64 * 128 * 144),
    ('irfft2 16x16, 64 ch + bias + GroupSort', 'irfft2_kernel<16,fm>', 64 * 144 * 8 + 64 * 256 * 4, 0.6e6),
    ('conv 256->64 @8x8 (s2d 1x1) + GroupSort', 'gemm_tc_kernel<s2d,64>', 64 * 256 * 4 + 64 * 64 * 4, 2 * 256 * 64 * 64),
    ('linear 4096->512 + GroupSort', 'gemm_tc_kernel<row,128>', 4096 * 4 + 512 * 4, 2 * 4096 * 512),
    ('linear 512->512 + GroupSort', 'gemm_tc_kernel<row,128>', 512 * 4 * 2, 2 * 512 * 512),
    ('fc 512->10', 'gemm_nt_kernel', 512 * 4 + 40, 2 * 512 * 10),
]


# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture (profiles/r01h_ncu_full_summary.md, one 4096-image launch)
# (profiles/r01c_ncu_full_summary.md; re-measure when a kernel changes).  Keyed by the kernel named in STAGES / the solver.
NCU_TRAFFIC_BYTES = {
    'rfft2_kernel<32>': 50.37e6 + 4.59e6,
    'irfft32_mix3_wide_kernel': 53.95e6 + 476.89e6,
    'gemm_tc_kernel<s2d,32>': 536.93e6 + 114.31e6,
    'rfft2_kernel<16,fm>': 134.24e6 + 100.68e6,
    'gemm_tc_kernel<row,128> batched': 156.84e6 + 245.81e6,
    'irfft2_kernel<16,fm>': 302.05e6 + 228.97e6,
    'gemm_tc_kernel<s2d,64>': 268.56e6 + 46.21e6,
    'dopri5_simplex_batched_kernel': 0.55e6,
}


def stage_roofline(ivp, x_dev, G, ts, pk, iters=10, cluster_size=0):
    """Device time of every kernel stage of one step (backbone over G*256 images, then the G-solve dopri5 launch),
    CUDA events on the launching stream, averaged over ``iters`` runs on rotating (> L2) inputs; the `roofline`
    object describes the kernel with the largest share of the step (launch time x launches per step)."""
    import ctypes
    from fiode_b200 import ops
    from fiode_b200._lib import lib
    L = lib()
    n_img = G * B
    pm = ivp.init_coordinates.param_map
    bw, keep = pm.backbone_weights()
    n_stage = min(n_img, L.fiode_backbone_slab_images())   # the stage events describe ONE slab of the backbone
    xs = [x.reshape(n_img, 3, 32, 32)[:n_stage] for x in x_dev]
    acc = [0.0] * len(STAGES)
    buf = (ctypes.c_float * 16)()
    L.fiode_backbone_enable_stage_timing(1)
    try:
        for i in range(iters + 2):
            ops.backbone_fwd(bw, xs[i % len(xs)])
            rc = L.fiode_backbone_stage_times(buf, 16)
            if rc == 0 and i >= 2:
                for j in range(len(STAGES)):
                    acc[j] += buf[j] / iters
    finally:
        L.fiode_backbone_enable_stage_timing(0)
    s = ops.backbone_fwd(bw, x_dev[0].reshape(n_img, 3, 32, 32))
    dw = ivp.dyn_fun.dyn_weights()
    h0 = torch.full((n_img, 10), 0.1, device=s.device)
    t_dev = ts.to(device=s.device, dtype=torch.float64)
    res = ops.dopri5_simplex_batched(dw, h0, s, G, t_dev, TOL, TOL, cluster_size=cluster_size)
    st = res.stats()[0]
    ms_solve = timed(lambda: ops.dopri5_simplex_batched(dw, h0, s, G, t_dev, TOL, TOL, cluster_size=cluster_size), iters) / iters
    n_slabs = -(-n_img // n_stage)              # backbone launches per step (one per slab)
    ridge = pk['bf16'] * 1e12 / 3 / (pk['hbm'] * 1e9)      # flop/byte above which the 3-MMA emulation is tensor-bound
    stages = []
    for (name, kern, by, fl), ms in zip(STAGES, acc):
        d = dict(stage=name, kernel=kern, ms=ms, launches_per_step=n_slabs, ms_per_step=ms * n_slabs)
        if by:
            d.update(gbs=by * n_stage / (ms * 1e-3) / 1e9 if ms > 0 else None, tflops=fl * n_stage / (ms * 1e-3) / 1e12 if ms > 0 else None,
                     bound='tensor' if fl / by > ridge else 'hbm')
            d['frac'] = (d['tflops'] / pk['bf16']) if d['bound'] == 'tensor' else (d['gbs'] / pk['hbm'])
        stages.append(d)
    solve_flop = sum(x.nfe for x in res.stats()) * RHS_FLOP * B
    stages.append(dict(stage=f'dopri5, {G} solves x {B} rows (cluster per solve)', kernel='dopri5_simplex_batched_kernel',
                       ms=ms_solve, launches_per_step=1, ms_per_step=ms_solve, tflops=solve_flop / (ms_solve * 1e-3) / 1e12, bound='latency',
                       note='%d dependent RHS evaluations per solve, %d rows per CTA: FP32 FFMA, latency-bound' % (st.nfe, B // max(cluster_size, 1) if cluster_size else 32)))
    ms_bb = sum(acc)
    bb_bytes = sum(x[2] for x in STAGES) * n_stage
    top = max((d for d in stages if 'tflops' in d or 'gbs' in d), key=lambda d: d['ms_per_step'])    # dominant = largest share of the step
    if top.get('bound') == 'hbm':
        rf = dict(bound='hbm', achieved=top['gbs'], peak=pk['hbm'], unit='GB/s', frac=top['gbs'] / pk['hbm'])
    else:
        rf = dict(bound='tensor', achieved=top['tflops'], peak=pk['bf16'], unit='TFLOP/s', frac=top['tflops'] / pk['bf16'])
    traffic = NCU_TRAFFIC_BYTES.get(top['kernel'])
    if traffic and top['kernel'] != 'dopri5_simplex_batched_kernel':
        traffic *= n_stage / 4096.             # the capture ran one 4096-image launch; these kernels stream, bytes scale with images
    rf.update(traffic=traffic, traffic_source='profiles/r01h_ncu_full_summary.md (4096-image launch, scaled to the %d images of this launch)' % n_stage,
              kernel=top['kernel'], stage=top['stage'], ms_per_launch=top['ms'], launches_per_step=top['launches_per_step'], peak_source=pk['source'],
              note='algorithmic bytes / flops of SURVEY 8(d) over the CUDA-event time of that launch; tensor-core stages run the '
                   'fp16x2-split FP32 emulation (3 MMAs per product), so their ceiling is peak / 3; see DESIGN.md section 4',
              parts=dict(backbone_ms=ms_bb, backbone_images=n_stage, backbone_algorithmic_gbs=bb_bytes / (ms_bb * 1e-3) / 1e9,
                         backbone_hbm_frac=bb_bytes / (ms_bb * 1e-3) / 1e9 / pk['hbm'],
                         backbone_tflops=BACKBONE_FLOP * n_stage / (ms_bb * 1e-3) / 1e12, dopri5_ms=ms_solve,
                         nfe=st.nfe, n_accept=st.n_accept, n_reject=st.n_reject, stages=stages))
    return rf


def run_b200(args):
    import torch.distributed as dist
    import fiode_b200
    from fiode_b200 import ops
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    assert torch.cuda.is_available(), 'bench.py needs a CUDA device (there is no CPU fallback)'
    torch.cuda.set_device(local)
    device = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=device)
    # pinned host buffers on the GPU's own NUMA node: bind this process to the device's CPU set while they are
    # allocated and while the GPU is driven; the CPU baseline later gets all cores back
    all_cpus = os.sched_getaffinity(0)
    numa = None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByUUID('GPU-' + str(torch.cuda.get_device_properties(local).uuid))
        pynvml.nvmlDeviceSetCpuAffinity(h)
        numa = len(os.sched_getaffinity(0))
    except Exception as e:           # affinity is an optimisation only
        numa = repr(e)
    pk = peaks()
    CS = args.cluster_size
    G = args.solves_per_step
    if G <= 0:
        # one wave of the cluster-per-solve launch: every solve beyond the co-resident count waits a full solve latency
        G = ops.dopri5_max_concurrent_solves(B, CS)
        G = G - G % 16 if G >= 16 else G          # a multiple of 16 batches (4096 images)
    ivp = build_model(device)
    ts = torch.linspace(0., T_MAX, 2)
    int_params = dict(method='dopri5', rtol=TOL, atol=TOL, options=dict(check=False, cluster_size=CS))

    # inputs: NBUF rotating device-resident sets, together larger than the 126 MB L2
    per_set = G * B * 3 * 32 * 32 * 4
    NBUF = max(2, int(math.ceil(160e6 / per_set)))
    gen = torch.Generator().manual_seed(1234 + rank)
    x_host = [torch.rand(G, B, 3, 32, 32, generator=gen).pin_memory() for _ in range(min(NBUF, 2))]
    x_dev = [x_host[i % len(x_host)].to(device) + 0 for i in range(NBUF)]
    y_dev = torch.empty(G, B, 10, device=device)
    y_host = torch.empty(G, B, 10).pin_memory()

    def step(xs, out):
        # G independent batch-256 solves: backbone once over G*256 images, one cluster-per-solve dopri5 launch
        out.copy_(ivp.forward_batches(xs, ts=ts, int_params=int_params))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (value) ------------------------------------------------------------
    for i in range(args.warmup):
        step(x_dev[i % NBUF], y_dev)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    fiode_b200.reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(x_dev[i % NBUF], y_dev)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    launches = fiode_b200.launch_count()
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms], device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * G * args.steps / (ms_max * 1e-3)

    # ---- end to end through the public API with host buffers (e2e) ------------------------------
    # Every step copies its own inputs host -> device (pinned) and its result device -> host inside the timed region.
    # The copy of step i+1 runs on a second stream while step i computes (two device buffers), which is how a
    # caller feeding batches from a loader would drive the API; each step still ends with its result on the host.
    copy_stream = torch.cuda.Stream(device=device)
    main_stream = torch.cuda.current_stream(device)
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]

    def e2e_measure(hosts, depth=1):
        """whole-job solves/s with `hosts` (pinned host tensors) as each step's input.  depth 1: the caller blocks on every
        step's result before submitting the next step; depth 2: it submits step i + 1, then collects the result of step i
        (what a loader-fed inference loop does), so the GPU does not idle while the host turns around."""
        xe = [torch.empty(hosts[0].shape, dtype=hosts[0].dtype, device=device) for _ in range(2)]
        yd = [torch.empty_like(y_dev) for _ in range(2)]
        yh = [torch.empty_like(y_host).pin_memory() for _ in range(2)]
        done = [torch.cuda.Event() for _ in range(2)]

        def prefetch(i):
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[i % 2])
                xe[i % 2].copy_(hosts[i % len(hosts)], non_blocking=True)
                ready[i % 2].record(copy_stream)

        def e2e_run(n):
            for k in range(2):
                consumed[k].record(main_stream)
            prefetch(0)
            for i in range(n):
                if i + 1 < n:
                    prefetch(i + 1)
                main_stream.wait_event(ready[i % 2])
                step(xe[i % 2], yd[i % 2])
                consumed[i % 2].record(main_stream)
                yh[i % 2].copy_(yd[i % 2], non_blocking=True)
                done[i % 2].record(main_stream)
                if depth == 1:
                    done[i % 2].synchronize()              # the caller needs this step's result on the host
                elif i >= 1:
                    done[(i - 1) % 2].synchronize()        # result of the previous step, while this one runs
            done[(n - 1) % 2].synchronize()

        e2e_run(max(2, args.warmup // 2))
        barrier()
        e0.record()
        e2e_run(args.steps)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return world * G * args.steps / (float(t.item()) * 1e-3)

    e2e_value = e2e_measure(x_host, depth=2)
    # the same call on the dataset's own 8-bit pixels (ToTensor's / 255 runs inside the first kernel): a quarter of the bytes
    x_host_u8 = [torch.randint(0, 256, (G, B, 3, 32, 32), dtype=torch.uint8, generator=gen).pin_memory() for _ in range(2)]
    e2e_u8_value = e2e_measure(x_host_u8, depth=2)
    e2e_u8_blocking = e2e_measure(x_host_u8, depth=1)

    cert = None
    if not args.no_certify:
        try:
            cert = certify_bench(ivp, device, pk, world)
        except Exception as e:       # secondary metric must not take the headline down
            cert = dict(error=repr(e))
    fi = None
    if not args.no_fi:
        try:
            fi = fi_step_bench(device, pk, world)
        except Exception as e:           # secondary metric must not take the headline down
            fi = dict(error=repr(e))
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    os.sched_setaffinity(0, all_cpus)
    # ---- per-kernel device times (CUDA events on the launching stream) and the roofline of the dominant kernel ----
    roofline = stage_roofline(ivp, x_dev, G, ts, pk, cluster_size=CS)

    # ---- CPU baseline beside it (bounded sample) ----------------------------------------------------
    cpu = None
    if not args.no_cpu_baseline and world == 1:          # the contract asks for it on rank 0 at N=1 only
        threads = os.cpu_count() or 1
        n_cpu = 100                      # about 12 s of host work at ~8 solves/s
        rate, secs = cpu_solve_rate(cpu_state_dict(ivp), n_cpu, threads)
        cpu = dict(value=rate, unit=UNIT, cores=threads, kind='port',
                   sample=f'{n_cpu} batch-256 solves (backbone + dopri5) in {secs:.1f} s on the same weights')
        # and the answer check that makes the timing meaningful
        from oracle import dynamics as od
        xc = x_host[0][0]
        want = od.ivp_forward(xc, cpu_state_dict(ivp), dict(alpha_1=100., alpha_2=20., sigma_1=0.02, scale_nominal=False),
                              t_max=T_MAX, tol=TOL)
        got = ivp(xc.to(device), ts=ts, int_params=int_params).cpu()
        cpu['max_abs_diff_vs_cuda'] = float((got - want).abs().max())

    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup,
                ms_per_step=ms_max / args.steps, higher_is_better=True, scaling='weak', vs_baseline=None,
                dtype='f32', data='synthetic',
                config=dict(workload='cifar_cayley_node_inference_dopri5_b256_tol1e-3_tmax1', batch=B,
                            solves_per_step=G, dopri5_cluster_size=CS, images_per_step=G * B, rtol=TOL, atol=TOL, t_max=T_MAX,
                            weights='random init, seed 0', l2='inputs rotate over %d device buffers (%.0f MB > L2)' % (
                                NBUF, NBUF * per_set / 1e6), parallelism=f'independent batches x{world}', host_cpus_bound_to_gpu_numa=numa),
                images_per_s=value * B,
                # e2e: the dataset's 8-bit pixels in pinned host memory -> result on the host; ToTensor (/255, dataset_loaders.py:76-78)
                # runs inside the first kernel, bit-identical to the float call (tests/test_gpu_parity.py).  e2e_f32_images: the
                # same loop fed with ToTensor's float32 output (4x the bytes; host-memory bound once 8 ranks copy at once).
                e2e=dict(value=e2e_u8_value, unit=UNIT, h2d_bytes_per_step=per_set // 4, d2h_bytes_per_step=G * B * 10 * 4,
                         input='uint8 pixels (G,256,3,32,32) in pinned host memory, ToTensor fused into the first kernel',
                         results_in_flight=2, blocking_value=e2e_u8_blocking,
                         note='every step: H2D copy of its inputs, the public forward_batches call, D2H copy of its result; the host '
                              'collects the result of step i after submitting step i + 1 (blocking_value: it waits for each result '
                              'before submitting the next step)'),
                e2e_f32_images=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=per_set, d2h_bytes_per_step=G * B * 10 * 4,
                                    input='float32 images in pinned host memory (the argument of model.forward)', results_in_flight=2),
                gpu_launches=launches, clocks=clocks, roofline=roofline, cpu_baseline=cpu, certify=cert, fi_step=fi)
    print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--solves-per-step', type=int, default=0, help='0 = as many as one dopri5 launch keeps co-resident')
    ap.add_argument('--cluster-size', type=int, default=4, help='CTAs per solve in the cluster-per-solve dopri5 launch')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-certify', action='store_true')
    ap.add_argument('--no-fi', action='store_true')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
