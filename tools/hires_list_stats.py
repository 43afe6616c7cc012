import sys, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from _harness import run_ours
from gaussian_gan_decoder_b200 import synthetic
dev = torch.device('cuda:0')
s = synthetic.raster_scene(2_000_000, size=2048, fov_deg=12.0, seed_offset=50)
s = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in s.items()}
o = run_ours(s, dev)
r = o['ranges'].view(-1, 2).long()
n = (r[:, 1] - r[:, 0])
print('R', o['R'], 'tiles', n.numel(), 'nonempty', int((n > 0).sum()), 'max', int(n.max()), 'mean nonempty', float(n[n > 0].float().mean()))
for a, b in ((1, 4096), (4097, 12288), (12289, 24576), (24577, 49152), (49153, 10**9)):
    m = (n >= a) & (n <= b)
    print(f'{a:6d}..{b:<10d} lists {int(m.sum()):6d} instances {int(n[m].sum()):11d}')
