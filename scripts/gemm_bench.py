"""Stand-alone timing of the tcgen05 GEMM core against cuBLAS (torch.matmul) on the head's shapes."""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from unified_video_action_b200 import _lib

dev = torch.device("cuda", 0)
L = _lib.lib()

def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters

for (M, N, K) in [(65536, 3072, 1024), (65536, 1024, 1024), (32768, 3072, 1024), (32768, 1024, 1024), (8192, 8192, 8192)]:
    A = torch.randn(M, K, device=dev).bfloat16()
    B = (torch.randn(N, K, device=dev) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device=dev)
    D = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    fl = 2.0 * M * N * K
    st = torch.cuda.current_stream().cuda_stream
    res = {}
    res["cublas"] = timeit(lambda: torch.matmul(A, B.t(), out=D))
    for name, lay in (("pair store", 24), ("quad store", 32), ("pair silu", 25), ("quad silu", 33)):
        def f():
            rc = L.uva_gemm_bf16_selftest(A.data_ptr(), B.data_ptr(), bias.data_ptr(), D.data_ptr(), M, N, K, 256, lay, st)
            assert rc == 0, L.uva_last_error()
        res[name] = timeit(f)
    print(f"M={M} N={N} K={K}: " + "  ".join(f"{k} {v*1e3:.1f}us {fl/v/1e9:.0f}TF" for k, v in res.items()), flush=True)
