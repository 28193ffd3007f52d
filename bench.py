#!/usr/bin/env python
"""bench.py — diffusion-head token-steps/s (BASELINE.json metric) on N B200s of one node.

Workload (config.workload): BASELINE config 3, "UVA PushT-M multitask inference, 100-step DDPM + CFG
over all masked tokens, batch 64 per GPU": DiffLoss.sample on 64 x 1024 cond rows || 64 x 1024 uncond
rows = 131 072 rows per GPU, C=16, cfg=3.0, temperature 0.95, clip_denoised=False, 100 denoise steps.
One bench "step" = one such sample() call = 13 107 200 row-steps per GPU.  token-steps/s is a
sampling metric, so the sampling config it is quoted on is the default workload; the other
BASELINE configs are reported as extra keys of the same JSON line at N=1 ("extras": C1 p50
predict_action head latency, C2 training rows/s) or selected with --workload:

    --workload train     C2: DiffLoss fwd+bwd N=16384 + DiffActLoss fwd+bwd N=256 per GPU, rows/s
    --workload latency   C1: DiffActLoss.sample B=1 (N=16) and B=56 (N=896), p50 over 200 calls
    --workload sweep     C5: DiffLoss.sample cfg=1, N = 1k .. 256k rows

    python bench.py --gpus 1 --steps 3 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # CPU arm: the oracle port of the reference head

value  : whole-job row-steps/s with z / start noise / step noise already resident in HBM, timed with
         CUDA events around K library calls (max over ranks).
e2e    : same metric through the public API DiffLoss.sample(z) with z in pinned HOST memory: the
         H2D copy of z, the RNG draws, the 100-step loop and the D2H copy of the samples are all timed.
roofline: dominant kernel = the AdaLN-modulation GEMM with the LayerNorm-modulate epilogue
         (umma_gemm2_kernel<256, EPI_LNMOD>; [rows,1024] x [1024,20480] per denoise step in 7 launches,
         61% of the algorithmic FLOPs), timed live with CUDA event pairs on its launch stream in a
         separate pass of the same workload with the two-lane overlap off (so each launch runs alone).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

C, Z, W, DEPTH, T_GEN = 16, 768, 1024, 6, 100
CFG, TEMP = 3.0, 0.95
FLOP_PER_ROW_STEP = 2 * (C * W + DEPTH * 5 * W * W + 2 * W * W + W * 2 * C)  # 67 207 168
ADA_FLOP_PER_ROW = 2 * (3 * DEPTH + 2) * W * W  # AdaLN modulation GEMM, per row per denoise step
# dram__bytes_read.sum + dram__bytes_write.sum of one EPI_LNMOD launch (65536 rows x 3072 columns) from the
# ncu --set full capture in profiles/ (None until captured for the current kernel)
NCU_TRAFFIC_BYTES = 653.3e6  # profiles/r02_ncu_full_summary.txt: 413.2 MB read + 240.1 MB written (r01: 654.9e6)
TRAIN_FLOP_PER_ROW = 6 * (C * W + 256 * W + W * W + Z * W + DEPTH * 5 * W * W + 2 * W * W + W * 2 * C)


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["bf16_tflops"]), float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured"
    return 1590.0, 1400.0, "fallback"


def _hbm_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"])
    return 6650.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons, pw = [], 0.0, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                pw.append(float(r[2]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        pw.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm), "power_w": pw[len(pw) // 2] if pw else None}


def make_head(dev, cls_name="DiffLoss"):
    """Reference init (DL:237-259) then the zero-initialised tensors re-drawn N(0, 0.02), seed 0, so
    the net is not identically zero (SURVEY §8d synthetic inputs)."""
    import torch

    from unified_video_action_b200 import DiffLoss

    torch.manual_seed(0)
    head = DiffLoss(C, Z, DEPTH, W, str(T_GEN), n_frames=4, language_emb_model=None, language_emb_model_type=None)
    g = torch.Generator().manual_seed(0)
    with torch.no_grad():
        for n, p in head.named_parameters():
            if "adaLN_modulation" in n or n.startswith("net.final_layer.linear"):
                p.copy_(0.02 * torch.randn(p.shape, generator=g))
    return head.to(dev)


def cpu_sample_rate(rows, steps, seed=0):
    """The oracle port of the reference head (fp32, torch CPU, all host threads) on `rows` rows x
    `steps` denoise steps of the same CFG workload -> row-steps/s."""
    import torch

    from oracle import uva_oracle as O

    P = O.to_torch_params(O.make_params(C, Z, W, DEPTH, seed=1))
    S = O.create_schedule(str(T_GEN))
    g = torch.Generator().manual_seed(seed)
    z = torch.randn(rows, Z, generator=g)
    half = torch.randn(rows // 2, C, generator=g)
    x0 = torch.cat([half, half])
    nz = torch.randn(steps, rows, C, generator=g)
    t0 = time.perf_counter()
    O.sample_loop(P, S, z, x0, nz, temperature=TEMP, cfg=CFG, clip_denoised=False, num_steps=steps)
    dt = time.perf_counter() - t0
    return rows * steps / dt, dt


def make_act_head(dev, A=2):
    """DiffActLoss (PushT: A=2; DAL:12-107) with the zero-initialised tensors re-drawn like make_head."""
    import torch

    from unified_video_action_b200 import DiffActLoss

    torch.manual_seed(1)
    head = DiffActLoss(A, Z, DEPTH, W, str(T_GEN), n_frames=4, language_emb_model=None, language_emb_model_type=None)
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for n, p in head.named_parameters():
            if "adaLN_modulation" in n or n.startswith("net.final_layer.linear"):
                p.copy_(0.02 * torch.randn(p.shape, generator=g))
    return head.to(dev)


def _p50(xs):
    xs = sorted(xs)
    return xs[len(xs) // 2]


def measure_latency(dev, batches=(1, 56), calls=200):
    """C1: p50 wall-clock latency of DiffActLoss.sample(z[B,1024,768]) (pre-net + RNG draws + 100-step
    DDPM loop, clip_denoised=True, cfg=1) with z on the device and the actions read back to the host,
    i.e. the head's share of predict_action (MAR:1037-1041)."""
    import torch

    head = make_act_head(dev)
    out = {}
    for B in batches:
        z = torch.randn(B, 1024, Z, device=dev)
        for _ in range(5):
            head.sample(z, temperature=TEMP, cfg=1.0).cpu()
        ts = []
        l0 = head.net.head().launch_count()
        for _ in range(calls):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            head.sample(z, temperature=TEMP, cfg=1.0).cpu()
            ts.append((time.perf_counter() - t0) * 1e3)
        launches = (head.net.head().launch_count() - l0) / calls
        ts.sort()
        out[f"B{B}_rows{16 * B}"] = {"p50_ms": ts[len(ts) // 2], "p90_ms": ts[int(len(ts) * 0.9)], "calls": calls,
                                    "row_steps_per_s": 16 * B * T_GEN / (ts[len(ts) // 2] / 1e3),
                                    "launches_per_call": launches}
        # production mode: step noise drawn in the kernel (Philox4x32-10) instead of 100 torch.randn calls that
        # keep the reference's RNG stream
        head.net.device_noise = True
        for _ in range(5):
            head.sample(z, temperature=TEMP, cfg=1.0).cpu()
        ts = []
        for _ in range(calls):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            head.sample(z, temperature=TEMP, cfg=1.0).cpu()
            ts.append((time.perf_counter() - t0) * 1e3)
        head.net.device_noise = None
        ts.sort()
        e = out[f"B{B}_rows{16 * B}"]
        e["p50_ms_device_noise"] = ts[len(ts) // 2]
        # latency-regime roofline: every denoise step has to stream the bf16 weights once
        wbytes = 2 * (2 * W + DEPTH * 5 * W * W + 2 * W * W + W * 4)  # input_proj, 6 x (AdaLN 3W^2 + MLP 2W^2), final AdaLN 2W^2, final linear
        e["us_per_step"] = e["p50_ms"] * 1e3 / T_GEN
        e["weight_gbs"] = wbytes * T_GEN / (e["p50_ms"] / 1e3) / 1e9
        e["weight_floor_us_per_step"] = wbytes / (_hbm_gbs() * 1e9) * 1e6
    return out


def measure_allreduce(dev, world, iters=10):
    """The path's only collective, alone: the two heads' flat gradient buffers (35.75 M + 35.71 M fp32 values, what
    the fused backward hands out) all-reduced through the C ABI, fp32 on the wire and bf16 on the wire.  Returns ms per
    step (both heads) and the bus bandwidth 2 (N-1)/N * bytes / time (max over ranks is taken by the caller)."""
    import torch

    from unified_video_action_b200 import _lib

    comm = _lib.GradComm()
    flats = [torch.randn(n, device=dev) for n in (35749920, 35706884)]
    out = {}
    for name, bf16 in (("fp32", False), ("bf16", True)):
        for f in flats:
            comm.all_reduce_(f, True, bf16=bf16)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            for f in flats:
                comm.all_reduce_(f, True, bf16=bf16)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        wire = sum(f.numel() for f in flats) * (2 if bf16 else 4)
        out[name] = {"ms": ms, "wire_bytes": wire, "bus_gbs": 2 * (world - 1) / world * wire / (ms / 1e3) / 1e9}
    return out


def measure_train(dev, world, B=16, iters=5, warmup=2, bf16_wire=False):
    """C2: one joint training step of the two heads (full_dynamic_model case): DiffLoss fwd+bwd on
    B x 1024 rows + DiffActLoss fwd+bwd on B x 16 rows, bf16 GEMMs, gradients wrt z and every head
    parameter; with world > 1 the head gradients are all-reduced over NCCL.  Returns rows/s
    per GPU and ms per step (CUDA events)."""
    import torch
    import torch.distributed as dist

    from unified_video_action_b200 import allreduce_head_grads

    vid, act = make_head(dev), make_act_head(dev)
    act.overlap_with(vid)  # what patch() sets up inside a MAR: the action head runs beside the video head
    if world > 1 and os.environ.get("UVA_EARLY_ALLREDUCE", "1") != "0":
        from unified_video_action_b200 import overlap_allreduce

        overlap_allreduce([vid, act], bf16=bf16_wire, in_backward=os.environ.get("UVA_IN_BACKWARD", "0") != "0")  # collectives start while the backward runs
    g = torch.Generator(device=dev).manual_seed(4321)
    z = torch.randn(B, 1024, Z, generator=g, device=dev, requires_grad=True)
    target = 0.2325 * torch.randn(B, 1024, C, generator=g, device=dev)
    mask = (torch.rand(B, 1024, generator=g, device=dev) < 0.85).float()
    actions = torch.rand(B, 16, 2, generator=g, device=dev) * 2 - 1

    def step():
        for m in (vid, act):
            for p in m.parameters():
                p.grad = None
        z.grad = None
        # the reference trains under Accelerate's bf16 autocast (workspace/train_unified_video_action_workspace.py:
        # mixed_precision), which is what the PyTorch pre-net of DiffActLoss (Conv2d 768->768 3x3 + Linears) sees;
        # the heads themselves take fp32 tensors and compute in bf16 either way.  UVA_BENCH_AUTOCAST=0: fp32 pre-net.
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=os.environ.get("UVA_BENCH_AUTOCAST", "1") != "0"):
            loss = vid(target, z, mask) + act(actions, z)
        loss.backward()
        if world > 1:
            allreduce_head_grads([vid, act], bf16=bf16_wire)
        return loss

    # at least 30 untimed steps: the small head's fwd / bwd replay from CUDA graphs that the library captures the second
    # time it sees a set of buffer pointers, and the caching allocator cycles through a few sets before it settles; a
    # capture inside the timed region is a single 20 - 170 ms step (seen up to the ~15th step of a fresh process).
    # p50_ms / max_ms below show the spread of the timed steps
    for _ in range(max(warmup, 30)):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    import gc

    gc.collect()
    gc.disable()  # a cyclic-GC pass of the host interpreter inside the timed region showed up as single 20 ms steps
    e0.record()
    for i in range(iters):
        loss = step()
        marks[i].record()
    e1.record()
    torch.cuda.synchronize()
    gc.enable()
    ms = e0.elapsed_time(e1) / iters  # the reported number: whole timed region / steps
    per = sorted(a.elapsed_time(b) for a, b in zip([e0] + marks[:-1], marks))  # spread of the single steps
    rows = B * 1024 + B * 16
    # optimizer + EMA + operand re-pack of both heads: fused AdamW/EMA pass, then load_weights
    import copy

    from unified_video_action_b200 import HeadAdamW

    opt = HeadAdamW([vid, act], lr=1e-4, betas=(0.9, 0.95), weight_decay=0.0, ema_modules=[copy.deepcopy(vid), copy.deepcopy(act)])
    step()
    opt.step(); vid.net.head(); act.net.head()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        opt.step()
        vid.net.head()
        act.net.head()
    e1.record()
    torch.cuda.synchronize()
    opt_ms = e0.elapsed_time(e1) / iters
    return {"ms_per_step": ms, "p50_ms": per[len(per) // 2], "max_ms": per[-1], "optimizer_ema_repack_ms": opt_ms, "rows_per_gpu": rows, "rows_per_s_per_gpu": rows / (ms / 1e3),
            "tflops_per_gpu": (B * 1024) * TRAIN_FLOP_PER_ROW / (ms / 1e3) / 1e12, "loss": float(loss.detach()),
            "collective": ("none (1 GPU)" if world == 1 else
                           f"NCCL all-reduce of the head gradients only (2 flat buffers, {'bf16' if bf16_wire else 'fp32'} on the wire), started from each head's backward"),
            "what": f"DiffLoss fwd+bwd {B}x1024 rows + DiffActLoss fwd+bwd {B}x16 rows (incl. its conv_fc pre-net, fused in the library), bf16"}


def _max_over_ranks(x, dev, world):
    import torch
    import torch.distributed as dist

    t = torch.tensor([float(x)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def train_extras(dev, world):
    """extras.C2_train at every N: the joint training step WITH the head-gradient all-reduce (the path's only
    collective), fp32 and bf16 on the wire, plus the collective alone with its bus bandwidth.
    Times are the max over ranks."""
    res = measure_train(dev, world, iters=20)
    res["ms_per_step"] = _max_over_ranks(res["ms_per_step"], dev, world)
    res["rows_per_s_per_gpu"] = res["rows_per_gpu"] / (res["ms_per_step"] / 1e3)
    res["tflops_per_gpu"] = (res["rows_per_gpu"] - 16 * 16) * TRAIN_FLOP_PER_ROW / (res["ms_per_step"] / 1e3) / 1e12
    if world > 1:
        b = measure_train(dev, world, iters=20, bf16_wire=True)
        res["bf16_wire"] = {"ms_per_step": _max_over_ranks(b["ms_per_step"], dev, world), "p50_ms": b["p50_ms"]}
        ar = measure_allreduce(dev, world)
        for k in ar:
            ar[k]["ms"] = _max_over_ranks(ar[k]["ms"], dev, world)
            ar[k]["bus_gbs"] = 2 * (world - 1) / world * ar[k]["wire_bytes"] / (ar[k]["ms"] / 1e3) / 1e9
        res["allreduce_alone"] = ar
    return res


def measure_gpu_eager(dev):
    """Like-for-like GPU baseline: the reference head's op graph executed by stock
    PyTorch eager kernels ON THE SAME B200 - the oracle port, which is the reference's torch code restated op for op
    (the reference package itself cannot travel to the GPU box) - in fp32 with TF32 off (how the reference samples)
    and under bf16 autocast.  C3-shaped: 8192 rows (cond || uncond, CFG 3.0) x 10 denoise steps; C1: 16 rows x 100
    steps, clip_denoised.  Baseline measurement only: none of this is on the product path."""
    import torch

    from oracle import uva_oracle as O

    P = {k: v.to(dev) for k, v in O.to_torch_params(O.make_params(C, Z, W, DEPTH, seed=1)).items()}
    S = O.create_schedule(str(T_GEN))
    out = {}
    g = torch.Generator(device=dev).manual_seed(5)
    for tag, rows, steps, cfg, clip in (("C3_8192rows_10steps", 8192, 10, CFG, False), ("C1_16rows_100steps", 16, 100, 1.0, True)):
        z = torch.randn(rows, Z, generator=g, device=dev)
        if cfg != 1.0:
            h = torch.randn(rows // 2, C, generator=g, device=dev)
            x0 = torch.cat([h, h])
        else:
            x0 = torch.randn(rows, C, generator=g, device=dev)
        nz = torch.randn(steps, rows, C, generator=g, device=dev)
        for name, ctx in (("fp32", torch.autocast("cuda", enabled=False)), ("bf16_autocast", torch.autocast("cuda", dtype=torch.bfloat16))):
            with ctx:
                O.sample_loop(P, S, z, x0, nz, temperature=TEMP, cfg=cfg, clip_denoised=clip, num_steps=min(steps, 3))
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                O.sample_loop(P, S, z, x0, nz, temperature=TEMP, cfg=cfg, clip_denoised=clip, num_steps=steps)
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
            out[f"{tag}_{name}"] = {"row_steps_per_s": rows * steps / dt, "ms": dt * 1e3, "us_per_step": dt * 1e6 / steps}
    out["what"] = "oracle port (the reference's torch op graph) run by PyTorch eager on this GPU; torch.backends.cuda.matmul.allow_tf32 = %s" % torch.backends.cuda.matmul.allow_tf32
    return out


def measure_libero(dev, B=32, calls=3):
    """C4 (Libero10 / multi-view, MAR:1037-1041, 1101-1103, 1120-1122): one sample_tokens iteration =
    diffactloss.sample (A = 10, clip_denoised, z [B,1024,768]) + diffloss.sample + diffloss_wrist.sample on the same
    B x 1024 rows, cfg 1 (shipped configs), B = 32.  ms per iteration and row-steps/s over all three heads."""
    import torch

    vid, wrist, act = make_head(dev), make_head(dev), make_act_head(dev, A=10)
    z = torch.randn(B, 1024, Z, device=dev)
    rows = z.reshape(B * 1024, Z)

    def it():
        a = act.sample(z, TEMP, cfg=1.0)
        v = vid.sample(rows, TEMP, 1.0)
        w = wrist.sample(rows, TEMP, 1.0)
        return a, v, w

    it()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(calls):
        a, v, w = it()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / calls
    n = B * 16 + 2 * B * 1024
    return {"ms_per_iteration": ms, "rows": n, "row_steps_per_s": n * T_GEN / (ms / 1e3), "B": B,
            "step_tflops": (B * 16 * 2 * (10 * W + DEPTH * 5 * W * W + 2 * W * W + W * 20) + 2 * B * 1024 * FLOP_PER_ROW_STEP) * T_GEN / (ms / 1e3) / 1e12,
            "what": "DiffActLoss.sample A=10 (B x 16 rows) + DiffLoss.sample + wrist DiffLoss.sample (B x 1024 rows each), 100 steps, cfg 1"}


def measure_sweep(dev, sizes=(1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072, 262144), dsteps=T_GEN):
    """C5: DiffLoss.sample, cfg=1, N = 1k .. 256k rows: row-steps/s per size (device-resident inputs)."""
    import torch

    lib_head = make_head(dev).net.head()
    out = {}
    for n in sizes:
        g = torch.Generator(device=dev).manual_seed(n)
        z = torch.randn(n, Z, generator=g, device=dev)
        x0 = torch.randn(n, C, generator=g, device=dev)
        nz = torch.randn(dsteps, n, C, generator=g, device=dev)
        f = lambda: lib_head.sample(z, x0, nz, temperature=TEMP, cfg=1.0, clip_denoised=False, k_begin=0, k_end=dsteps)
        f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 3 if n <= 32768 else 1
        e0.record()
        for _ in range(reps):
            f()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        out[str(n)] = {"ms": ms, "row_steps_per_s": n * dsteps / (ms / 1e3)}
        del z, x0, nz
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port: the reference
    is Python and cannot travel to the GPU box), all host threads, bounded sample per step."""
    import torch

    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    try:  # torchrun exports OMP_NUM_THREADS=1: the CPU arm uses every host core this process may run on
        torch.set_num_threads(len(os.sched_getaffinity(0)))
    except Exception:
        pass
    cores = torch.get_num_threads()
    rows, dsteps = 2048, 20
    for _ in range(args.warmup):
        cpu_sample_rate(256, 1)
    rates, times = [], []
    for k in range(args.steps):
        r, dt = cpu_sample_rate(rows, dsteps, seed=k)
        rates.append(r)
        times.append(dt)
    total = rows * dsteps * args.steps / sum(times)
    sample = f"{rows} rows ({rows // 2} cond || {rows // 2} uncond) x {dsteps} of the 100 denoise steps per bench step, fp32 torch CPU"
    line = {
        "impl": "reference", "metric": "diffusion-head token-steps/sec", "value": total, "unit": "row-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C3: DiffLoss.sample, 100-step DDPM + CFG 3.0, C=16 (bounded CPU sample)", "rows_per_step": rows,
                   "denoise_steps_per_step": dsteps},
        "cpu_baseline": {"value": total, "unit": "row-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": total, "unit": "row-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="sample", choices=["sample", "train", "latency", "sweep", "libero", "eager"])
    ap.add_argument("--rows", type=int, default=131072, help="rows per GPU (cond || uncond)")
    ap.add_argument("--denoise-steps", type=int, default=T_GEN)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the C1 latency / C2 training extras of the default line")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    os.environ["UVA_B200_PRECISION"] = args.precision

    if args.workload != "sample":
        if args.workload in ("libero", "eager"):
            line = None
            if rank == 0:
                if args.workload == "libero":
                    line = {"metric": "diffusion-head token-steps/sec", "unit": "row-steps/s", "n_gpus": 1, "higher_is_better": True,
                            "dtype": args.precision, "data": "synthetic", "config": {"workload": "C4: Libero10 multi-view sample_tokens iteration, B=32"},
                            "detail": measure_libero(dev, calls=max(args.steps, 1))}
                    line["value"] = line["detail"]["row_steps_per_s"]
                else:
                    line = {"metric": "diffusion-head token-steps/sec", "unit": "row-steps/s", "n_gpus": 1, "higher_is_better": True,
                            "dtype": "f32 / bf16 autocast", "data": "synthetic", "impl": "pytorch-eager-on-gpu (baseline)",
                            "config": {"workload": "reference op graph by PyTorch eager on this GPU"}, "detail": measure_gpu_eager(dev)}
            if line is not None:
                print(json.dumps(line), flush=True)
            if world > 1:
                dist.destroy_process_group()
            return
        if args.workload == "train":
            res = measure_train(dev, world, iters=max(args.steps, 1), warmup=max(args.warmup, 1),
                                bf16_wire=os.environ.get("UVA_BF16_WIRE", "0") != "0")
            t = torch.tensor([res["ms_per_step"]], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res["ms_per_step"] = float(t.item())
            line = {"metric": "diffusion-head training rows/sec (fwd+bwd)", "unit": "rows/s", "n_gpus": world,
                    "value": world * res["rows_per_gpu"] / (res["ms_per_step"] / 1e3), "ms_per_step": res["ms_per_step"],
                    "higher_is_better": True, "scaling": "weak", "dtype": args.precision, "data": "synthetic",
                    "config": {"workload": "C2: " + res["what"], "collective": res["collective"]},
                    "detail": res}
        elif args.workload == "latency":
            line = {"metric": "p50 predict_action head latency", "unit": "ms", "n_gpus": 1, "higher_is_better": False,
                    "dtype": args.precision, "data": "synthetic",
                    "config": {"workload": "C1: DiffActLoss.sample, z [B,1024,768] on device, A=2, 100 steps, clip_denoised"},
                    "detail": measure_latency(dev)} if rank == 0 else None
        else:
            line = {"metric": "diffusion-head token-steps/sec", "unit": "row-steps/s", "n_gpus": 1, "higher_is_better": True,
                    "dtype": args.precision, "data": "synthetic", "config": {"workload": "C5: DiffLoss.sample cfg=1, N sweep"},
                    "detail": measure_sweep(dev, dsteps=args.denoise_steps)} if rank == 0 else None
        if rank == 0 and line is not None:
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    head = make_head(dev)
    lib_head = head.net.head()
    rows, K, Wm, dsteps = args.rows, args.steps, max(args.warmup, 0), args.denoise_steps

    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    z = torch.randn(rows, Z, generator=g, device=dev)
    halfn = torch.randn(rows // 2, C, generator=g, device=dev)
    x0 = torch.cat([halfn, halfn])
    step_noise = torch.randn(dsteps, rows, C, generator=g, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def call(k_end=dsteps):
        return lib_head.sample(z, x0, step_noise, temperature=TEMP, cfg=CFG, clip_denoised=False, k_begin=0, k_end=k_end)

    for _ in range(Wm):
        call()
    barrier()
    l0 = lib_head.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        barrier()
        e0.record()
        for _ in range(K):
            out = call()
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    launches = lib_head.launch_count() - l0
    assert torch.isfinite(out).all(), "non-finite samples"
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * rows * dsteps * K / (ms / 1e3)

    # ---- roofline pass: the dominant kernel timed alone (event pairs around every launch, one lane) ----
    # The kernel is timed ALONE and compared with the BURST peak, which the driver measured as the best of 10 matmuls on an
    # idle GPU: give the clocks the same chance (the timed region above leaves the chip at the power cap's ~1.1 GHz), then
    # time a short pass; `achieved` stays the AVERAGE over all timed launches of the pass
    time.sleep(1.5)
    rsteps = min(dsteps, 4)
    lib_head.timing_enable(True)
    call(rsteps)
    torch.cuda.synchronize()
    lib_head.timing_enable(False)
    ada_ms, ada_n = lib_head.timing_read(0)
    cat_ms = {0: ada_ms}
    for cat in (1, 2, 3, 4):
        cat_ms[cat], _ = lib_head.timing_read(cat)

    # ---- e2e through the public API, host buffers ----
    z_host = z.cpu().pin_memory()
    out_host = torch.empty(rows, C, dtype=torch.float32).pin_memory()
    z_dev = torch.empty_like(z)

    def call_e2e():
        z_dev.copy_(z_host, non_blocking=True)
        o = head.sample(z_dev, temperature=TEMP, cfg=CFG)
        out_host.copy_(o, non_blocking=True)

    e2e = None
    if dsteps == T_GEN:
        del step_noise
        call_e2e()
        barrier()
        e0.record()
        for _ in range(K):
            call_e2e()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": world * rows * T_GEN * K / (float(t.item()) / 1e3), "unit": "row-steps/s",
               "h2d_bytes_per_step": rows * Z * 4, "d2h_bytes_per_step": rows * C * 4}

    if rank == 0:
        burst, sustained, src = _peaks()
        ada_tflops = (ADA_FLOP_PER_ROW * rows * rsteps / (ada_ms / 1e3) / 1e12) if ada_ms > 0 else None
        line = {
            "metric": "diffusion-head token-steps/sec", "value": value, "unit": "row-steps/s", "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": {"workload": "C3: DiffLoss.sample 100-step DDPM + CFG 3.0, B=64/GPU -> 131072 rows (cond||uncond), C=16",
                       "rows_per_gpu": rows, "denoise_steps": dsteps, "cfg": CFG, "temperature": TEMP,
                       "l2": "inputs larger than L2 (z 403 MB, step noise 839 MB per call)", "parallelism": f"rows x{world}"},
            "e2e": e2e, "gpu_launches": int(launches),
            "roofline": {"bound": "tensor", "achieved": ada_tflops, "peak": burst, "unit": "TFLOP/s",
                         "frac": (ada_tflops / burst) if ada_tflops else None, "traffic": NCU_TRAFFIC_BYTES,
                         "traffic_note": "ncu dram bytes of one 65536-row x 3072-column launch; algorithmic bytes of that launch: 671 MB (ys + h read, a + gate written)",
                         "kernel": "umma_gemm2_kernel<256, EPI_LNMOD> (AdaLN modulation GEMM + LayerNorm-modulate epilogue)",
                         "launches_timed": int(ada_n), "avg_launch_ms": (ada_ms / ada_n) if ada_n else None,
                         "flop_per_launch_avg": ADA_FLOP_PER_ROW * rows * rsteps / ada_n if ada_n else None,
                         "share_of_gemm_time": ada_ms / sum(cat_ms.values()) if ada_ms > 0 else None,
                         "timed": f"{rsteps} denoise steps after 1.5 s idle, chunks one after the other, CUDA event pair per launch, average over the launches",
                         "peak_source": f"MEASURED_PEAKS.json bf16_tflops burst ({src}) because the kernel is timed alone in a short pass; sustained {sustained}"},
            "step_tflops": value / world * FLOP_PER_ROW_STEP / 1e12,
            "step_frac_of_burst_peak": value / world * FLOP_PER_ROW_STEP / 1e12 / burst,
            "clocks": clk.summary(),
        }
        if not args.no_cpu_baseline and world == 1:
            crow, csteps = 2048, 50
            try:
                torch.set_num_threads(len(os.sched_getaffinity(0)))
            except Exception:
                pass
            rate, dt = cpu_sample_rate(crow, csteps)
            line["cpu_baseline"] = {"value": rate, "unit": "row-steps/s", "cores": torch.get_num_threads(), "kind": "port",
                                    "sample": f"{crow} rows x {csteps} denoise steps of the same CFG workload ({dt:.1f} s), fp32 torch CPU oracle port"}
    extras_on = not args.no_extras and dsteps == T_GEN
    if extras_on:
        del z, x0, z_dev
        torch.cuda.empty_cache()
        # the training step with the path's only collective is measured at EVERY N (all ranks take part)
        c2 = train_extras(dev, world)
    if rank == 0:
        if extras_on:
            line["extras"] = {"C2_train": c2}
            if world == 1:
                line["extras"]["C1_latency"] = measure_latency(dev, calls=100)
                line["extras"]["C4_libero"] = measure_libero(dev)
                line["extras"]["C5_sweep"] = measure_sweep(dev, sizes=(1024, 2048, 4096, 8192, 16384, 32768, 65536))
                line["extras"]["gpu_eager_baseline"] = measure_gpu_eager(dev)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
