"""Is the C2 training step CPU- (launch-) or GPU-bound?  CPU enqueue time vs CUDA-event time per step."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
vid, act = bench.make_head(dev), bench.make_act_head(dev)
B = 16
g = torch.Generator(device=dev).manual_seed(4321)
z = torch.randn(B, 1024, 768, generator=g, device=dev, requires_grad=True)
target = 0.2325 * torch.randn(B, 1024, 16, generator=g, device=dev)
mask = (torch.rand(B, 1024, generator=g, device=dev) < 0.85).float()
actions = torch.rand(B, 16, 2, generator=g, device=dev) * 2 - 1

def step(which):
    for m in (vid, act):
        for p in m.parameters(): p.grad = None
    z.grad = None
    loss = 0
    if which in ("both", "video"): loss = loss + vid(target, z, mask)
    if which in ("both", "action"): loss = loss + act(actions, z)
    loss.backward()

for which in ("both", "video", "action"):
    for _ in range(3): step(which)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(10): step(which)
    e1.record(); t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"{which}: cpu enqueue {1e3*(t1-t0)/10:.2f} ms/step, gpu {e0.elapsed_time(e1)/10:.2f} ms/step", flush=True)
