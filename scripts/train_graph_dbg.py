import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
act = bench.make_act_head(dev)
g = torch.Generator(device=dev).manual_seed(1)
z = torch.randn(16, 1024, 768, generator=g, device=dev, requires_grad=True)
actions = torch.rand(16, 16, 2, generator=g, device=dev) * 2 - 1
for i in range(8):
    for p in act.parameters(): p.grad = None
    z.grad = None
    sys.stderr.write(f"--- step {i}\n")
    act(actions, z).backward()
torch.cuda.synchronize()
# host time split
import cProfile, pstats
def step():
    for p in act.parameters(): p.grad = None
    z.grad = None
    act(actions, z).backward()
for _ in range(3): step()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(20): step()
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(14); pstats.Stats(pr).sort_stats("tottime").print_stats(22)
