"""B=56 (896 rows) latency path: three library calls (eager, capture+replay, replay) for an ncu launch list."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
dev = torch.device("cuda", 0)
head = bench.make_act_head(dev)
N, C = 896, 2
g = torch.Generator(device=dev).manual_seed(1)
zz = torch.randn(N, 768, device=dev, generator=g)
noise = torch.randn(N, C, device=dev, generator=g)
sn = torch.randn(100, N, C, device=dev, generator=g)
lh = head.net.head()
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    lh.sample(zz, noise, sn, temperature=0.95, cfg=1.0, clip_denoised=True)
torch.cuda.synchronize()
