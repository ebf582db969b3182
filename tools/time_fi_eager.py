"""Dev tool: host and device time of one eager FI training step (forward / backward), after warm-up."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = [sys.argv[0]]
import runpy
g = runpy.run_path(os.path.join(os.path.dirname(__file__), 'prof_fi.py'))
mod, x, y, Bt, step = g['mod'], g['x'], g['y'], g['Bt'], g['step']
for _ in range(5):
    step()
torch.cuda.synchronize()
tf = tb = 0.0
N = 10
t_all0 = time.perf_counter()
for _ in range(N):
    for p in mod.parameters():
        p.grad = None
    t0 = time.perf_counter()
    loss = mod.compute_loss(x, y, Bt, act='relu')
    t1 = time.perf_counter()
    loss.backward()
    t2 = time.perf_counter()
    tf += t1 - t0; tb += t2 - t1
torch.cuda.synchronize()
t_all1 = time.perf_counter()
print(f'FIODE_SAVE_MINV={os.environ.get("FIODE_SAVE_MINV", "1")}: host forward {tf / N * 1e3:.2f} ms, host backward {tb / N * 1e3:.2f} ms, wall per step {(t_all1 - t_all0) / N * 1e3:.2f} ms')
