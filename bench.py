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
                note='forward + backward to all 33 parameters + one flat all-reduce, replayed from a CUDA graph; the Cayley '
                     'transforms of all layers are prepared up front on side streams (layers.prefetch_orthogonal), weight gradients '
                     'run on a branch beside the data gradients; FFMA GEMMs (gradients are not fp16-range safe for the split)')


# per image: algorithmic HBM bytes (fp32 activations in + out; spectra complex64) and flops of each backbone stage
#   name, kernel, bytes/img, flop/img, (kernel key, launch index) in profiles/ncu_traffic.json
STAGES = [
    ('weight packing (per call)', 'pack_w_kernel x5 + mix_real_weights', 0, 0, None),
    ('rfft2 32x32, 3 ch (+Normalize)', 'rfft2_kernel<32>', 3 * 1024 * 4 + 3 * 544 * 8, 0.12e6, ('rfft2_kernel<32, 0, 0>', 0)),
    ('mix 3->32 + inverse DFT 32x32 (tcgen05) + bias + GroupSort', 'irfft32_mix3_tc2_kernel', 3 * 544 * 8 + 32 * 1024 * 4, 0.42e6 + 6.7e6, ('irfft32_mix3_tc2_kernel', 0)),
    ('conv 128->32 @16x16 (s2d 1x1) + GroupSort', 'gemm_tc_kernel<s2d,32>', 32 * 1024 * 4 + 32 * 256 * 4, 2 * 128 * 32 * 256, ('gemm_tc_kernel<1, 32, 0>', 0)),
    ('rfft2 16x16, 32 ch', 'rfft2_kernel<16,fm>', 32 * 256 * 4 + 32 * 144 * 8, 0.3e6, ('rfft2_kernel<16, 1, 0>', 0)),
    ('mixing 144 x (64->128 real)', 'gemm_tc_kernel<row,128> batched', (32 + 64) * 144 * 8, 2 * 64 * 128 * 144, ('gemm_tc_kernel<0, 128, 1>', 0)),
    ('irfft2 16x16, 64 ch + bias + GroupSort', 'irfft2_kernel<16,fm>', 64 * 144 * 8 + 64 * 256 * 4, 0.6e6, ('irfft2_kernel<16, 2>', 0)),
    ('conv 256->64 @8x8 (s2d 1x1) + GroupSort', 'gemm_tc_kernel<s2d,64>', 64 * 256 * 4 + 64 * 64 * 4, 2 * 256 * 64 * 64, ('gemm_tc_kernel<1, 64, 0>', 0)),
    ('linear 4096->512 + GroupSort', 'gemm_tc_kernel<row,128>', 4096 * 4 + 512 * 4, 2 * 4096 * 512, ('gemm_tc_kernel<0, 128, 1>', 1)),
    ('linear 512->512 + GroupSort', 'gemm_tc_kernel<row,128>', 512 * 4 * 2, 2 * 512 * 512, ('gemm_tc_kernel<0, 128, 1>', 2)),
    ('fc 512->10', 'gemm_nt_kernel', 512 * 4 + 40, 2 * 512 * 10, ('gemm_nt_kernel', 0)),
]


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed `ncu --set full` capture of this code
    (profiles/ncu_traffic.json, written by `tools/ncu_summary.py traffic <capture>`; re-generated whenever a kernel changes --
    a number measured under the profiler cannot be taken inside this run)."""
    p = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')
    if not os.path.exists(p):
        return None
    return json.load(open(p))


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
    nt = ncu_traffic()
    for (name, kern, by, fl, nkey), ms in zip(STAGES, acc):
        d = dict(stage=name, kernel=kern, ms=ms, launches_per_step=n_slabs, ms_per_step=ms * n_slabs)
        if nt and nkey and nkey[0] in nt['kernels'] and len(nt['kernels'][nkey[0]]) > nkey[1]:
            # the capture ran one backbone pass over nt['backbone_images'] images; these kernels stream, bytes scale with images
            d['traffic'] = nt['kernels'][nkey[0]][nkey[1]]['dram_bytes'] * n_stage / nt['backbone_images']
            if by:
                d['traffic_over_algorithmic'] = d['traffic'] / (by * n_stage)
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
    bb_traffic = sum(d.get('traffic', 0.0) for d in stages[:len(STAGES)])
    top = max((d for d in stages if 'tflops' in d or 'gbs' in d), key=lambda d: d['ms_per_step'])    # dominant = largest share of the step
    if top.get('bound') == 'hbm':
        rf = dict(bound='hbm', achieved=top['gbs'], peak=pk['hbm'], unit='GB/s', frac=top['gbs'] / pk['hbm'])
    else:
        rf = dict(bound='tensor', achieved=top['tflops'], peak=pk['bf16'], unit='TFLOP/s', frac=top['tflops'] / pk['bf16'])
    traffic = top.get('traffic')
    if top['kernel'] == 'dopri5_simplex_batched_kernel' and nt and 'dopri5_simplex_kernel<1>' in nt['kernels']:
        traffic = nt['kernels']['dopri5_simplex_kernel<1>'][0]['dram_bytes']
    rf.update(traffic=traffic, traffic_source=('profiles/ncu_traffic.json <- %s (%d-image pass, scaled to the %d images of this launch)' % (
                  nt['source'], nt['backbone_images'], n_stage)) if nt else None,
              kernel=top['kernel'], stage=top['stage'], ms_per_launch=top['ms'], launches_per_step=top['launches_per_step'], peak_source=pk['source'],
              note='algorithmic bytes / flops of SURVEY 8(d) over the CUDA-event time of that launch; tensor-core stages run the '
                   'fp16x2-split FP32 emulation (3 MMAs per product), so their ceiling is peak / 3; see DESIGN.md section 4',
              parts=dict(backbone_ms=ms_bb, backbone_images=n_stage, backbone_algorithmic_gbs=bb_bytes / (ms_bb * 1e-3) / 1e9,
                         backbone_hbm_frac=bb_bytes / (ms_bb * 1e-3) / 1e9 / pk['hbm'],
                         backbone_ncu_traffic_bytes=bb_traffic, backbone_minimal_bytes_per_image=61e3,
                         backbone_tflops=BACKBONE_FLOP * n_stage / (ms_bb * 1e-3) / 1e12, dopri5_ms=ms_solve,
                         nfe=st.nfe, n_accept=st.n_accept, n_reject=st.n_reject, stages=stages))
    return rf


def segway_bench(device, pk, world, iters=10):
    """BASELINE config 1: Segway NN-controller forward-invariance grid check (control/certify_segway.py:35-59): the
    300 x 53 x 300 = 4.77 M-point grid (r = 0.01) generated from its index, shell mask V in [lb, ub], barrier derivative
    of the closed loop at the shell points, ordered compaction.  Grid slabs shard across ranks (weak scaling would need
    a finer grid; here every rank evaluates the WHOLE grid, N replicas)."""
    from fiode_b200.control import segway
    torch.manual_seed(0)
    ctrl = torch.nn.Sequential(torch.nn.Linear(3, 32), torch.nn.ReLU(), torch.nn.Linear(32, 32), torch.nn.ReLU(), torch.nn.Linear(32, 1))
    sd = {'model.' + k: v.detach() for k, v in ctrl.state_dict().items()}
    P = torch.eye(3) + 0.1 * torch.randn(3, 3, generator=torch.Generator().manual_seed(2))
    sizes, r = [math.pi / 12, 1.5, 1.5], [0.01] * 3
    n = 1
    for a in segway.grid_axes(sizes, r):
        n *= len(a)
    fn = lambda: segway.certify_grid(sd, P, sizes, r, level=0.15, device=device)
    eta, val, vmax = fn()
    ms = timed(fn, iters) / iters
    t = torch.tensor([ms], device=device)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    flop = 2 * (3 * 32 + 32 * 32 + 32) + 60                  # controller MLP at a shell point + closed-form dynamics
    out = dict(metric='Segway grid pts/s (certify_segway shell + barrier derivative)', value=world * n / (ms * 1e-3), unit='grid pts/s (whole job)',
               workload=f'grid {n} pts (r=0.01), NNController(3,1,32) seed 0, P = I + 0.1 randn seed 2, level 0.15; every rank the whole grid',
               ms_per_call=ms, shell_points=int(eta.shape[0]), max_barrier_derivative=vmax,
               note='public call segway.certify_grid: allocation of the (n, 3) candidate buffer, ONE kernel (index -> point -> shell test -> '
                    'controller + dynamics at shell points -> ordered compaction) and the 8-byte status read; the grid is generated, never read: '
                    '0 B in, 16 B out per shell point',
               roofline=dict(bound='issue', achieved=n * 12 / (ms * 1e-3) / 1e9, unit='G lane-instr/s (about 12 FP32 instr per grid point for the shell test)',
                             shell_tflops=eta.shape[0] * flop / (ms * 1e-3) / 1e12, ffma_peak_tflops=FFMA_PEAK_TFLOPS))
    return out, (sd, P, sizes, r)


def cpu_legs(ivp, device, seg_cfg, threads):
    """CPU baselines beside the secondary metrics (SURVEY 8d 'CPU baseline beside it'), oracle ports on the host cores,
    each a bounded sample: certify = 1 image x 1 of the 10 chunks of grid_40 (extrapolated, labelled), FI step = the oracle's
    backbone + fi_loss forward and autograd backward on B=128 x S=256, Segway = the whole 4.77 M-point grid."""
    from oracle import certify as oc, dynamics as od, layers as ol, segway as oseg, sampler as osm
    from tests import fixtures as fx
    from fiode_b200 import grid as fgrid
    torch.set_num_threads(threads)
    sd = cpu_state_dict(ivp)
    out = {}
    # certify_lipschitz, one image, one chunk (certify_lipschitz.py:104-143)
    try:
        T = 40
        gz = fgrid.boundary_grid_numerators(10, T, device=device)
        n_all = gz.shape[0]
        chunk = n_all // 10
        z = gz[:chunk].cpu().numpy()
        label = 3
        grid = oc.grid_for_label(z, T, label)
        cfg = dict(alpha_1=100., alpha_2=20., sigma_1=0.02, scale_nominal=False)
        s = torch.randn(1, 10, generator=torch.Generator().manual_seed(1236)) * 0.5
        Q = od.dyn_Q(sd, 'model.dyn_fun.')
        t0 = time.perf_counter()
        oc.certify_image(grid, s, label, sd, cfg, eps=0.141, T=T, batches=1, Q=Q)
        dt = time.perf_counter() - t0
        out['certify'] = dict(value=chunk / dt, unit='grid pts/s', cores=threads, kind='port',
                              sample=f'1 image x 1 chunk of {chunk} grid_40 points in {dt:.1f} s (oracle certify_image); full config '
                                     f'(10 000 images x {n_all} points) extrapolates to {10000 * n_all / (chunk / dt) / 3600:.0f} h')
    except Exception as e:
        out['certify'] = dict(error=repr(e))
    # FI training step (pl_modules.py:390-506): forward + backward to the 33 parameters
    try:
        Bt, S = 128, 256
        cfg = dict(alpha_1=100., alpha_2=20., sigma_1=0.02, scale_nominal=True)
        x = fx.synth_images(Bt, seed=1234)
        y = fx.synth_labels(Bt, seed=1235)
        e = torch.empty(S, 10).exponential_(generator=torch.Generator().manual_seed(5))
        h = osm.uniform_simplex(e).repeat(Bt, 1)
        params = {k: v.clone().requires_grad_(v.is_floating_point() and ('weight' in k or 'bias' in k or 'alpha' in k)) for k, v in sd.items()}

        def fi_step():
            s_ = ol.backbone_forward(x, params, 'model.init_coordinates.param_map.')
            loss = od.fi_loss(h, s_.repeat_interleave(S, 0), y.repeat_interleave(S), params, cfg, 2.0, prefix='model.dyn_fun.')[0]
            loss.backward()
            return float(loss.detach())
        fi_step()
        ts_ = []
        for _ in range(2):
            t0 = time.perf_counter()
            fi_step()
            ts_.append(time.perf_counter() - t0)
        dt = sorted(ts_)[0]
        out['fi_step'] = dict(value=1.0 / dt, unit='steps/s', cores=threads, kind='port',
                              sample=f'B=128 x S=256 oracle backbone + fi_loss forward + autograd backward, best of 2 after a warm-up: {dt:.2f} s per step')
    except Exception as e:
        out['fi_step'] = dict(error=repr(e))
    try:
        sdc, P, sizes, r = seg_cfg
        t0 = time.perf_counter()
        eta, val, mask = oseg.certify_grid(sdc, P, sizes, r, level=0.15)
        dt = time.perf_counter() - t0
        out['segway'] = dict(value=mask.numel() / dt, unit='grid pts/s', cores=threads, kind='port',
                             sample=f'whole {mask.numel()}-point grid once (numpy meshgrid + torch shell mask + controller at {eta.shape[0]} shell points): {dt:.2f} s')
    except Exception as e:
        out['segway'] = dict(error=repr(e))
    return out


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
    # the public call with its DEFAULT options: the end-of-solve status read (torchdiffeq's assertions) is on
    # graph=True: the pass is replayed from a CUDA graph (fiode_b200/models.py), so the host turn-around after the status read is one
    # graph launch instead of ~15 kernel launches; value_eager is the same call without it
    int_params = dict(method='dopri5', rtol=TOL, atol=TOL, options=dict(cluster_size=CS, graph=not args.no_graph))
    int_params_eager = dict(method='dopri5', rtol=TOL, atol=TOL, options=dict(cluster_size=CS))
    int_params_nocheck = dict(method='dopri5', rtol=TOL, atol=TOL, options=dict(check=False, cluster_size=CS))

    # inputs: NBUF rotating device-resident sets, together larger than the 126 MB L2
    per_set = G * B * 3 * 32 * 32 * 4
    NBUF = max(2, int(math.ceil(160e6 / per_set)))
    gen = torch.Generator().manual_seed(1234 + rank)
    x_host = [torch.rand(G, B, 3, 32, 32, generator=gen).pin_memory() for _ in range(min(NBUF, 2))]
    x_dev = [x_host[i % len(x_host)].to(device) + 0 for i in range(NBUF)]
    y_dev = torch.empty(G, B, 10, device=device)
    y_host = torch.empty(G, B, 10).pin_memory()

    def step(xs, out, ip=int_params):
        # G independent batch-256 solves: backbone once over G*256 images, one cluster-per-solve dopri5 launch
        out.copy_(ivp.forward_batches(xs, ts=ts, int_params=ip))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed_blocks(run_k_steps, min_seconds=args.min_seconds, max_blocks=40):
        """Blocks of EXACTLY args.steps steps, each bracketed by barrier + synchronize and timed with CUDA events (max over
        ranks), repeated until >= min_seconds have been timed; returns the per-block times in ms."""
        blocks = []
        while True:
            barrier()
            e0.record()
            run_k_steps(args.steps)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device=device)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            blocks.append(float(t.item()))
            if sum(blocks) >= min_seconds * 1e3 or len(blocks) >= max_blocks:
                return blocks

    def rate(blocks):
        med = sorted(blocks)[len(blocks) // 2]
        return world * G * args.steps / (med * 1e-3), med

    # ---- device-resident timing (value) ------------------------------------------------------------
    for i in range(max(args.warmup, 3)):
        step(x_dev[i % NBUF], y_dev)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    fiode_b200.reset_launch_count()
    step(x_dev[0], y_dev, int_params_eager)          # the pass's own kernels, counted on one eager pass (graph replays launch the same nodes)
    launches_per_step = fiode_b200.launch_count()
    blocks = timed_blocks(lambda n: [step(x_dev[i % NBUF], y_dev) for i in range(n)])
    launches = launches_per_step * args.steps
    clocks = sampler.stop() if rank == 0 else None
    value, ms_block = rate(blocks)
    blocks_nc = timed_blocks(lambda n: [step(x_dev[i % NBUF], y_dev, int_params_nocheck) for i in range(n)])
    value_nc, ms_block_nc = rate(blocks_nc)
    value_eager = rate(timed_blocks(lambda n: [step(x_dev[i % NBUF], y_dev, int_params_eager) for i in range(n)]))[0]

    # ---- end to end through the public API with host buffers (e2e) ------------------------------
    # Every step copies its own inputs host -> device (pinned) and its result device -> host inside the timed region.
    # The copy of step i+1 runs on a second stream while step i computes (two device buffers), which is how a
    # caller feeding batches from a loader would drive the API; each step still ends with its result on the host.
    copy_stream = torch.cuda.Stream(device=device)
    main_stream = torch.cuda.current_stream(device)
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]

    def e2e_measure(hosts, depth=1, ip=int_params):
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
                step(xe[i % 2], yd[i % 2], ip)
                consumed[i % 2].record(main_stream)
                yh[i % 2].copy_(yd[i % 2], non_blocking=True)
                done[i % 2].record(main_stream)
                if depth == 1:
                    done[i % 2].synchronize()              # the caller needs this step's result on the host
                elif i >= 1:
                    done[(i - 1) % 2].synchronize()        # result of the previous step, while this one runs
            done[(n - 1) % 2].synchronize()

        e2e_run(max(2, args.warmup // 2))
        return rate(timed_blocks(e2e_run))[0]

    # e2e: float32 images (the argument of the reference's model.forward), default options (status check on)
    e2e_value = e2e_measure(x_host, depth=2)
    e2e_blocking = e2e_measure(x_host, depth=1)
    e2e_nocheck = e2e_measure(x_host, depth=2, ip=int_params_nocheck)
    # the same call on the dataset's own 8-bit pixels (ToTensor's / 255 runs inside the first kernel): a quarter of the bytes
    x_host_u8 = [torch.randint(0, 256, (G, B, 3, 32, 32), dtype=torch.uint8, generator=gen).pin_memory() for _ in range(2)]
    e2e_u8_value = e2e_measure(x_host_u8, depth=2)
    e2e_u8_nocheck = e2e_measure(x_host_u8, depth=2, ip=int_params_nocheck)

    # ---- the literal drop-in seam: module(x) on ONE batch of 256 (models.py:203-242), default options, host buffers ----
    def single_call(graph=False):
        xh = x_host[0][0]
        xd = torch.empty(B, 3, 32, 32, device=device)
        yh = torch.empty(B, 10).pin_memory()
        n = max(50, args.steps)
        ip = dict(method='dopri5', rtol=TOL, atol=TOL, options=dict(graph=True)) if graph else dict(method='dopri5', rtol=TOL, atol=TOL)

        def run(k):
            for _ in range(k):
                xd.copy_(xh, non_blocking=True)
                yh.copy_(ivp(xd, ts=ts, int_params=ip), non_blocking=True)
                torch.cuda.synchronize()
        run(10)
        barrier()
        t0 = time.perf_counter()
        e0.record()
        run(n)
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        t = torch.tensor([max(e0.elapsed_time(e1), wall * 1e3)], device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item()) / n
        return dict(value=world / (ms * 1e-3), unit=UNIT, ms_per_call=ms, calls=n, h2d_bytes_per_call=B * 3 * 32 * 32 * 4, d2h_bytes_per_call=B * 40,
                    note='IVP.forward(x) on one float32 batch of 256 from pinned host memory, default int_params (status check on), result copied '
                         'to the host and waited for before the next call: one solve in flight, latency-bound (backbone on 256 images + one '
                         'cooperative dopri5 launch); forward_batches is the same arithmetic with G solves in flight')
    try:
        e2e_single = single_call()
        g1 = single_call(graph=True)
        e2e_single['with_options_graph'] = dict(value=g1['value'], ms_per_call=g1['ms_per_call'],
                                                note='the same call with options=dict(graph=True): backbone + cooperative solve replayed from a CUDA graph')
    except Exception as e:
        e2e_single = dict(error=repr(e))

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
    seg, seg_cfg = None, None
    if not args.no_segway:
        try:
            seg, seg_cfg = segway_bench(device, pk, world)
        except Exception as e:
            seg = dict(error=repr(e))
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    os.sched_setaffinity(0, all_cpus)
    # ---- per-kernel device times (CUDA events on the launching stream) and the roofline of the dominant kernel ----
    roofline = stage_roofline(ivp, x_dev, G, ts, pk, cluster_size=CS)

    # ---- BASELINE config 5: dopri5 batch / tolerance sweep with the oracle's step counts beside the GPU's (summary; the
    #      table is tools/sweep.py -> profiles/) ------------------------------------------------------------------
    sweep = None
    if not args.no_sweep and world == 1:
        try:
            sys.path.insert(0, os.path.join(ROOT, 'tools'))
            import sweep as sweep_tool
            rows = sweep_tool.sweep(iters=3, oracle_path=os.path.join(ROOT, 'profiles', 'r02_sweep_oracle.json'), dev=device)
            sweep = sweep_tool.summary(rows)
            sweep['table'] = 'tools/sweep.py (B 64..8192 x tol 1e-2..1e-5 x 3 dynamics x cooperative / cluster launches); profiles/r02_sweep_dopri5.md'
        except Exception as e:
            sweep = dict(error=repr(e))

    # ---- CPU baseline beside it (bounded sample) ----------------------------------------------------
    cpu = None
    if not args.no_cpu_baseline and world == 1:          # the contract asks for it on rank 0 at N=1 only
        threads = os.cpu_count() or 1
        n_cpu = 100                      # about 12 s of host work at ~8 solves/s
        rate_cpu, secs = cpu_solve_rate(cpu_state_dict(ivp), n_cpu, threads)
        cpu = dict(value=rate_cpu, unit=UNIT, cores=threads, kind='port',
                   sample=f'{n_cpu} batch-256 solves (backbone + dopri5) in {secs:.1f} s on the same weights')
        # and the answer check that makes the timing meaningful
        from oracle import dynamics as od
        xc = x_host[0][0]
        want = od.ivp_forward(xc, cpu_state_dict(ivp), dict(alpha_1=100., alpha_2=20., sigma_1=0.02, scale_nominal=False),
                              t_max=T_MAX, tol=TOL)
        got = ivp(xc.to(device), ts=ts, int_params=int_params).cpu()
        cpu['max_abs_diff_vs_cuda'] = float((got - want).abs().max())
        legs = cpu_legs(ivp, device, seg_cfg, threads) if seg_cfg is not None else {}
        for name, obj in (('certify', cert), ('fi_step', fi), ('segway', seg)):
            if isinstance(obj, dict) and name in legs:
                obj['cpu_baseline'] = legs[name]

    d2h = G * B * 10 * 4
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                ms_per_step=ms_block / args.steps, higher_is_better=True, scaling='weak', vs_baseline=None,
                dtype='f32', data='synthetic',
                config=dict(workload='cifar_cayley_node_inference_dopri5_b256_tol1e-3_tmax1', batch=B,
                            solves_per_step=G, dopri5_cluster_size=CS, images_per_step=G * B, rtol=TOL, atol=TOL, t_max=T_MAX,
                            weights='random init, seed 0', l2='inputs rotate over %d device buffers (%.0f MB > L2)' % (
                                NBUF, NBUF * per_set / 1e6), parallelism=f'independent batches x{world}', host_cpus_bound_to_gpu_numa=numa,
                            options='end-of-solve status check on (the default: one 4-byte read + sync per step), options.graph=%s (CUDA-graph replay of the pass)' % (not args.no_graph)),
                timing=dict(blocks=len(blocks), steps_per_block=args.steps, timed_seconds=sum(blocks) / 1e3, block_ms_median=ms_block,
                            block_ms_min=min(blocks), block_ms_max=max(blocks),
                            note='value = the median block of EXACTLY --steps steps; blocks repeat until >= 0.5 s have been timed'),
                value_no_check=value_nc, ms_per_step_no_check=ms_block_nc / args.steps, value_eager_launches=value_eager,
                images_per_s=value * B,
                # e2e: float32 images (the reference's argument) in pinned host memory -> result on the host, default options
                e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=per_set, d2h_bytes_per_step=d2h,
                         input='float32 images (G,256,3,32,32) in pinned host memory (the argument of model.forward), default options (status check on)',
                         results_in_flight=2, blocking_value=e2e_blocking, value_no_check=e2e_nocheck,
                         note='every step: H2D copy of its inputs, the public forward_batches call, D2H copy of its result; the host '
                              'collects the result of step i after submitting step i + 1 (blocking_value: it waits for each result '
                              'before submitting the next step)'),
                # the dataset's 8-bit pixels: ToTensor (/255, dataset_loaders.py:76-78) runs inside the first kernel, bit-identical to the float
                # call (tests/test_gpu_parity.py), a quarter of the H2D bytes
                e2e_u8_pixels=dict(value=e2e_u8_value, unit=UNIT, h2d_bytes_per_step=per_set // 4, d2h_bytes_per_step=d2h, value_no_check=e2e_u8_nocheck,
                                   input='uint8 pixels in pinned host memory, ToTensor fused into the first kernel', results_in_flight=2),
                e2e_single_call=e2e_single,
                gpu_launches=launches, gpu_launches_per_step=launches_per_step, clocks=clocks, roofline=roofline, cpu_baseline=cpu, certify=cert, fi_step=fi, segway=seg, sweep=sweep)
    print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--solves-per-step', type=int, default=0, help='0 = as many as one dopri5 launch keeps co-resident')
    ap.add_argument('--cluster-size', type=int, default=4, help='CTAs per solve in the cluster-per-solve dopri5 launch')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-certify', action='store_true')
    ap.add_argument('--no-fi', action='store_true')
    ap.add_argument('--no-segway', action='store_true')
    ap.add_argument('--no-sweep', action='store_true')
    ap.add_argument('--no-graph', action='store_true', help='eager launches in the headline loop instead of options.graph=True')
    ap.add_argument('--min-seconds', type=float, default=0.5, help='repeat the K-step block until this much has been timed')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
