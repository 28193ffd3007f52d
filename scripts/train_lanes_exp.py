"""Does a two-lane training pass pay?  One head on 16384 rows vs two handles on 8192 rows each on two streams
(each handle's persistent GEMM grids limited to half of the SMs through set_train_reserve)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import uva_oracle as O
from unified_video_action_b200 import _lib

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
P = O.make_params(16, 768, 1024, 6, seed=1)
names = [n for n, _ in O.param_specs(16, 768, 1024, 6)]
params = [torch.from_numpy(P[n]).to(dev) for n in names]


def mk(reserve):
    h = _lib.Head(16, 768, 1024, 6, 1000, 100, _lib.MODE_BF16)
    h.load_weights(params)
    h.set_train_reserve(reserve)
    return h


N = 16384
g = torch.Generator(device=dev).manual_seed(0)
target = torch.randn(N, 16, generator=g, device=dev) * 0.5
z = torch.randn(N, 768, generator=g, device=dev)
mask = (torch.rand(N, generator=g, device=dev) < 0.85).float()
t = torch.randint(0, 1000, (N,), generator=g, device=dev)
noise = torch.randn(N, 16, generator=g, device=dev)
one = torch.ones(1, device=dev)


def grads():
    return [torch.zeros(P[n].shape, dtype=torch.float32, device=dev) for n in names]


def run(heads, streams, parts, iters=10):
    gs = [grads() for _ in heads]
    def step():
        ev = torch.cuda.current_stream().record_event()
        for h, s, (a, b), gr in zip(heads, streams, parts, gs):
            s.wait_event(ev)
            with torch.cuda.stream(s):
                h.loss_fwd(target[a:b], z[a:b], mask[a:b], t[a:b], noise[a:b])
                h.loss_bwd(one, target[a:b], z[a:b], mask[a:b], t[a:b], noise[a:b], gr)
        for s in streams:
            torch.cuda.current_stream().wait_stream(s)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


s0, s1 = torch.cuda.Stream(), torch.cuda.Stream()
for rep in range(2):
    print("one head, 16384 rows:            %.3f ms" % run([mk(0)], [s0], [(0, N)]))
    print("two heads x 8192 rows, 74 SMs each: %.3f ms" % run([mk(74), mk(74)], [s0, s1], [(0, N // 2), (N // 2, N)]))
    print("two heads x 8192 rows, all SMs:     %.3f ms" % run([mk(0), mk(0)], [s0, s1], [(0, N // 2), (N // 2, N)]))
    print("two heads x 8192 rows, one stream:  %.3f ms" % run([mk(0), mk(0)], [s0, s0], [(0, N // 2), (N // 2, N)]))
