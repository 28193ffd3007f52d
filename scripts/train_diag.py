"""Where does a C2 training step go: GPU time vs host enqueue time, per head (diagnostic, GPU box)."""
import sys, time, os
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


def run(which, iters=20):
    def step():
        for m in (vid, act):
            for p in m.parameters():
                p.grad = None
        z.grad = None
        loss = 0
        if "v" in which:
            loss = loss + vid(target, z, mask)
        if "a" in which:
            loss = loss + act(actions, z)
        loss.backward()
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(iters):
        step()
    t1 = time.perf_counter(); e1.record()
    torch.cuda.synchronize()
    print(f"{which:3s} overlap={os.environ.get('UVA_ACT_OVERLAP','1')}: gpu {e0.elapsed_time(e1)/iters:7.3f} ms/step, host enqueue {(t1-t0)/iters*1e3:7.3f} ms/step")


for w in ("v", "a", "va"):
    run(w)
os.environ["UVA_ACT_OVERLAP"] = "0"
run("va")
