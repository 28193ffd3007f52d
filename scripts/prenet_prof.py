"""What the DiffActLoss pre-net costs at B=1 (diagnostic)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from torch.profiler import profile, ProfilerActivity
dev = torch.device("cuda", 0)
head = bench.make_act_head(dev)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
z = torch.randn(B, 1024, 768, device=dev)
def wall(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    ts.sort(); return ts[len(ts) // 2]
with torch.no_grad():
    print("prenet default ms", wall(lambda: head.prenet(z)))
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3): head.prenet(z)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=12, max_name_column_width=70))
    torch.backends.cudnn.benchmark = True
    print("prenet cudnn.benchmark ms", wall(lambda: head.prenet(z)))
    with torch.autocast("cuda", dtype=torch.bfloat16):
        print("prenet bf16 autocast ms", wall(lambda: head.prenet(z)))
    torch.backends.cuda.matmul.allow_tf32 = True
    print("prenet tf32 matmul ms", wall(lambda: head.prenet(z)))
