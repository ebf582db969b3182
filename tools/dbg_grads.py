import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import fixtures as fx
from tests.test_gpu_fi import build_learning, dev
sd = fx.synth_state_dict()
for name in ['uniform_sn1', 'mixed_sn0', 'boundary_sn0']:
    c = fx.load_golden('fi.pt')[name]
    mod = build_learning(sd, c).train()
    x, y = dev(c['x']), dev(c['y'])
    var = [dev(v) if v is not None else None for v in c['variates']]
    loss = mod.compute_loss(x, y, c['B'], act='relu', variates=var)
    loss.backward()
    print(name, 'loss', float(loss), float(c['loss']))
    named = dict(mod.named_parameters())
    for k, wn in c['grad_norms'].items():
        g = named[k].grad
        line = f'  {k[6:]:55s} norm {float(g.norm()):.4e} ref {float(wn):.4e}'
        if k in c['grads']:
            w = c['grads'][k]
            if w.numel() > 1:
                line += f'  cos {float(torch.nn.functional.cosine_similarity(g.cpu().flatten(), w.flatten(), dim=0)):.4f}'
            else:
                line += f'  val {float(g):.4e} ref {float(w):.4e}'
        print(line)
