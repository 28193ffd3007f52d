"""Is the C2 step host-bound?  Host time to ENQUEUE a step vs device time per step."""
import os, sys, time, gc
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
vid, act = bench.make_head(dev), bench.make_act_head(dev)
act.overlap_with(vid)
B = 16
g = torch.Generator(device=dev).manual_seed(4321)
z = torch.randn(B, 1024, bench.Z, generator=g, device=dev, requires_grad=True)
target = 0.2325 * torch.randn(B, 1024, bench.C, generator=g, device=dev)
mask = (torch.rand(B, 1024, generator=g, device=dev) < 0.85).float()
actions = torch.rand(B, 16, 2, generator=g, device=dev) * 2 - 1


def step(which="both"):
    for m in (vid, act):
        for p in m.parameters():
            p.grad = None
    z.grad = None
    with torch.autocast("cuda", dtype=torch.bfloat16):
        if which == "both":
            loss = vid(target, z, mask) + act(actions, z)
        elif which == "vid":
            loss = vid(target, z, mask)
        else:
            loss = act(actions, z)
    loss.backward()


for which in ("both", "vid", "act"):
    for _ in range(40):
        step(which)
    torch.cuda.synchronize()
    gc.collect(); gc.disable()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 30
    e0.record()
    t0 = time.perf_counter()
    for _ in range(n):
        step(which)
    host = (time.perf_counter() - t0) / n * 1e3
    e1.record()
    torch.cuda.synchronize()
    gc.enable()
    print(f"{which:5s}: host enqueue {host:.3f} ms/step, device {e0.elapsed_time(e1) / n:.3f} ms/step")
