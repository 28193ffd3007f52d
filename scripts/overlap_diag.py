"""Do the two heads' training kernels overlap on the GPU when nothing but two streams is involved?
(library calls only: no autograd, no pre-net)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
vid, act = bench.make_head(dev), bench.make_act_head(dev)
hv, ha = vid.net.head(), act.net.head()
g = torch.Generator(device=dev).manual_seed(1)
def inputs(N, C):
    return (torch.randn(N, C, generator=g, device=dev), torch.randn(N, 768, generator=g, device=dev),
            torch.randint(0, 1000, (N,), generator=g, device=dev), torch.randn(N, C, generator=g, device=dev))
iv, ia = inputs(16384, 16), inputs(256, 2)
def grads(m):
    return [torch.zeros_like(p) for p in m.net.ordered_params()]
gv, ga = grads(vid), grads(act)
one = torch.ones(1, device=dev)
def run(h, i, gr):
    tgt, z, t, nz = i
    h.loss_fwd(tgt, z, None, t, nz)
    h.loss_bwd(one, tgt, z, None, t, nz, gr)
sa, sb = torch.cuda.Stream(), torch.cuda.Stream(priority=-1)
def timed(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / iters * 1e3
def only_v():
    with torch.cuda.stream(sa): run(hv, iv, gv)
def only_a():
    with torch.cuda.stream(sb): run(ha, ia, ga)
def both():
    with torch.cuda.stream(sa): run(hv, iv, gv)
    with torch.cuda.stream(sb): run(ha, ia, ga)
def both_a_first():
    with torch.cuda.stream(sb): run(ha, ia, ga)
    with torch.cuda.stream(sa): run(hv, iv, gv)
print("video only  %.3f ms" % timed(only_v))
print("action only %.3f ms" % timed(only_a))
print("both (v enqueued first) %.3f ms" % timed(both))
print("both (a enqueued first) %.3f ms" % timed(both_a_first))
