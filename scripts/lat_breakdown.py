"""Where the C1 latency goes: pre-net, RNG draws, library call (CUDA events + wall clock)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

dev = torch.device("cuda", 0)
head = bench.make_act_head(dev)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
z = torch.randn(B, 1024, 768, device=dev)
N, C = 16 * B, 2

def wall(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    ts.sort(); return ts[len(ts) // 2]

with torch.no_grad():
    print("prenet ms", wall(lambda: head.prenet(z)))
    zz = head.prenet(z).reshape(N, -1).contiguous()
    def draws():
        noise = torch.randn(N, C).to(dev)
        sn = torch.empty(100, N, C, device=dev)
        for k in range(100): torch.randn(N, C, out=sn[k])
        return noise, sn
    print("rng draws ms", wall(draws))
    noise, sn = draws()
    lh = head.net.head()
    print("lib sample ms", wall(lambda: lh.sample(zz, noise, sn, temperature=0.95, cfg=1.0, clip_denoised=True)))
    print("full sample ms", wall(lambda: head.sample(z, temperature=0.95, cfg=1.0).cpu()))
