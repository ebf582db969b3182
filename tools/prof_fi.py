"""Dev tool: one FI training step (B=128, S=256) under `ncu --profile-from-start off` or with torch's profiler off;
prints the CUDA-event time of the step's phases."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

dev = torch.device('cuda', 0)
torch.cuda.set_device(dev)
# build the module exactly like bench.fi_step_bench, but keep hold of it
import fiode_b200  # noqa: E402
from fiode_b200 import models, pl_modules, lya_cands  # noqa: E402
from fiode_b200.sampling import sampler as smp, sampler_schedulers as sch  # noqa: E402
from fiode_b200.dynamics.classification import OrthoClassDynProjectSimplexLips  # noqa: E402
from fiode_b200.dynamics.init_coordinates import UniformInitFun  # noqa: E402
from fiode_b200.dynamics.output_coordinates import DefaultOutputFun  # noqa: E402

Bt, S = 128, 256
torch.manual_seed(0)
pm = models.make_ortho_KWLarge_Concat(3, 10, mu=bench.MU, std=bench.STD, out_dim=10, act='GroupSort')
dyn = OrthoClassDynProjectSimplexLips(n_hidden=10, activation='ReLU', mlp_size=128, alpha_1=100., alpha_2=20., sigma_1=0.02,
                                      x_dim=10, cayley=True, dropout=0.5, scale_nominal=True, kappa=2.0, kappa_length=0)
mod = pl_modules.LyapunovLearning(
    order=1, h_sample_size=S, h_dist_lim=15, lya_cand=lya_cands.DecisionBoundary(on_simplex=True),
    sampler=smp.CompositeSampler((10,), [smp.UniformSimplexSampling((10,)), smp.CorrectConeSampling((10,))]),
    sampler_scheduler=sch.CompositeSamplerScheduler([sch.ConstantScheduler(0.5), sch.ConstantScheduler(0.5)], [1., 1.]),
    dynamics=dyn, output=DefaultOutputFun(), init_fun=UniformInitFun((10,), param_map=pm), t_max=1.0,
    train_ode_tol=1e-3, val_ode_tol=1e-3, epoch_off_scale=10).to(dev).train()
mod.current_epoch, mod.global_step = 0, 10 ** 9
x = torch.rand(Bt, 3, 32, 32, generator=torch.Generator().manual_seed(1234)).to(dev)
y = torch.randint(0, 10, (Bt,), generator=torch.Generator().manual_seed(1235)).to(dev)


def step():
    for p in mod.parameters():
        p.grad = None
    loss = mod.compute_loss(x, y, Bt, act='relu')
    loss.backward()
    return loss


for _ in range(3):
    step()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
ev[0].record()
loss = mod.compute_loss(x, y, Bt, act='relu')
ev[1].record()
loss.backward()
ev[2].record()
torch.cuda.synchronize()
print(f'forward {ev[0].elapsed_time(ev[1]):.2f} ms, backward {ev[1].elapsed_time(ev[2]):.2f} ms')
torch.cuda.profiler.start()
step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()

if len(sys.argv) > 1 and sys.argv[1] == 'graph':
    import traceback
    del loss                     # a live loss keeps last step's autograd graph (default-stream AccumulateGrad nodes) alive
    from fiode_b200.graphs import GraphedStep
    try:
        g = GraphedStep(step, warmup=2)
        g(); g()
        torch.cuda.synchronize()
        print('graphed step:', bench.timed(g, 20) / 20, 'ms; loss', float(g().detach()))
        print('eager step  :', bench.timed(step, 20) / 20, 'ms')
    except Exception:
        traceback.print_exc()

if len(sys.argv) > 2 and sys.argv[2] == 'trace':
    # kernel timeline of one graph replay (kineto): per-stream start / duration, for the critical path
    import json
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(4):          # back to back: the host-side graph launch of replay i+1 overlaps replay i, as in the timed loop
            g()
        torch.cuda.synchronize()
    os.makedirs('gpurun_out', exist_ok=True)
    prof.export_chrome_trace('gpurun_out/fi_trace_raw.json')
    ev = json.load(open('gpurun_out/fi_trace_raw.json'))['traceEvents']
    ks = [e for e in ev if e.get('cat') in ('kernel', 'gpu_memcpy', 'gpu_memset')]
    ks.sort(key=lambda e: e['ts'])
    per = len(ks) // 4
    print('replay starts (us):', [round(ks[i * per]['ts'] - ks[0]['ts']) for i in range(4)])
    ks = ks[3 * per:]
    t0 = ks[0]['ts']
    with open('gpurun_out/fi_trace.txt', 'w') as f:
        for e in ks:
            f.write(f"{e['ts'] - t0:9.1f} {e['dur']:8.1f} s{e['args'].get('stream', '?'):<4} {e['name'][:100]}\n")
    os.remove('gpurun_out/fi_trace_raw.json')
    print('trace:', len(ks), 'kernels, span', ks[-1]['ts'] + ks[-1]['dur'] - t0, 'us')
