"""Where the latency path (UVA_SMALL_ROWS) hands over to the chunked throughput path: ms per 100-step call."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
res = bench.measure_sweep(dev, sizes=tuple(int(b) for b in sys.argv[1].split(",")))
print("small_rows", os.environ.get("UVA_SMALL_ROWS", "default"), {k: round(v["ms"], 2) for k, v in res.items()})
