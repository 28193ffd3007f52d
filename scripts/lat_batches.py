"""C1 latency over small batches: where the chain kernel hands over to the per-step graph path (UVA_CHAIN_ROWS)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
res = bench.measure_latency(dev, batches=tuple(int(b) for b in sys.argv[1].split(",")), calls=40)
print(os.environ.get("UVA_CHAIN_ROWS", "default"), {k: (round(v["p50_ms"], 2), v["launches_per_call"]) for k, v in res.items()})
