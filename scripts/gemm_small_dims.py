"""Do the tcgen05 GEMM forms accept dimensions smaller than one TMA box (K = 32 < 64 in the data-gradient form,
M = 4 / 32 < 64 in the weight-gradient form)?  Needed by the final-layer backward on tensor cores."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from unified_video_action_b200 import _lib
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(0)
def rnd(*s): return torch.randn(*s, generator=g, device=dev).to(torch.bfloat16)
for layout, M, N, K in ((9, 4096, 1024, 32), (1, 100, 1024, 32), (2, 4, 1024, 4096), (2, 32, 1024, 16384), (10, 32, 1024, 4096)):
    if layout & 7 == 1:
        A, B = rnd(M, K), rnd(K, N); ref = A.float() @ B.float()
    else:
        A, B = rnd(K, M), rnd(K, N); ref = A.float().t() @ B.float()
    try:
        D = _lib.gemm_bf16_selftest(A, B, None, 256, layout)
        torch.cuda.synchronize()
        err = float((D - ref).abs().max() / ref.abs().max())
        print(layout, M, N, K, "rel err", err)
    except Exception as e:
        print(layout, M, N, K, "FAILED", str(e)[:200])
