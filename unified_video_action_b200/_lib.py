"""ctypes binding of ``libuva_head.so`` (C ABI in ``include/uva_head.h``).

PyTorch is used only as the owner of device memory and streams; every tensor crosses the
boundary as a raw device pointer.  There is no fallback: if the library cannot be built /
loaded, or no B200 is present, the calls raise.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import threading
from typing import Dict, List, Optional, Sequence

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_HERE, "libuva_head.so")
CSRC = os.path.join(_HERE, "csrc")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
]

MODE_FP32, MODE_BF16 = 0, 1


def _sources() -> List[str]:
    out = [os.path.join(_ROOT, "include", "uva_head.h")]
    for f in sorted(os.listdir(CSRC)):
        if f.endswith((".cu", ".cuh", ".hpp", ".inc")):
            out.append(os.path.join(CSRC, f))
    return out


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(s) > t for s in _sources())


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/uva_head.cu for sm_100a into the in-tree libuva_head.so (nvcc cross-compiles
    without a GPU)."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "nvcc")
    tmp = LIB_PATH + ".tmp.%d" % os.getpid()
    extra = os.environ.get("NVCC_EXTRA", "").split()  # e.g. -DUVA_DEBUG for the time-attribution switches
    cmd = [nvcc] + NVCC_FLAGS + extra + ["-o", tmp, os.path.join(CSRC, "uva_head.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    os.replace(tmp, LIB_PATH)
    if verbose:
        print(res.stderr)
    return LIB_PATH


class _Desc(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in ("C", "Z", "W", "depth", "T_train", "T_gen", "mode")]


class _Tail(ctypes.Structure):
    """UvaDecoderTail (include/uva_head.h)."""
    _fields_ = [("ln_w", ctypes.c_void_p), ("ln_b", ctypes.c_void_p), ("temporal", ctypes.c_void_p),
                ("spatial", ctypes.c_void_p), ("eps", ctypes.c_float), ("T", ctypes.c_int32), ("S", ctypes.c_int32),
                ("tokens_per_sample", ctypes.c_int64), ("skip", ctypes.c_int64)]


class DecoderTail:
    """Host-side description of the MAR decoder tail (mar_con_unified.py:702-724) for ``Head.sample_from_decoder``:
    ``decoder_norm`` weight / bias / eps, ``diffusion_temporal_embed`` [1,T,Z], ``diffusion_spatial_embed`` [1,S,Z],
    the number of leading text tokens dropped.  Keeps the fp32 device tensors alive."""

    def __init__(self, ln_weight, ln_bias, eps, temporal, spatial, skip: int = 0):
        self.keep = [_f32c(t.detach(), "decoder tail") for t in (ln_weight, ln_bias, temporal, spatial)]
        w, b, te, se = self.keep
        Z = w.numel()
        self.keep[2] = te = te.reshape(-1, Z)
        self.keep[3] = se = se.reshape(-1, Z)
        self.Z, self.T, self.S, self.skip, self.eps = Z, te.shape[0], se.shape[0], int(skip), float(eps)

    def struct(self, tokens_per_sample: int) -> "_Tail":
        w, b, te, se = self.keep
        return _Tail(w.data_ptr(), b.data_ptr(), te.data_ptr(), se.data_ptr(), self.eps, self.T, self.S,
                     int(tokens_per_sample), self.skip)


_lib = None
_lock = threading.Lock()

# every symbol include/uva_head.h declares (tests check that all of them are exported)
SYMBOLS = [
    "uva_head_num_params", "uva_head_param_numel", "uva_head_create", "uva_head_destroy",
    "uva_schedule_table", "uva_head_load_weights", "uva_head_workspace_bytes", "uva_head_net_forward",
    "uva_head_net_forward_static", "uva_head_static_max_rows",
    "uva_head_prenet_num_params", "uva_head_prenet_load_weights", "uva_head_prenet_workspace_bytes",
    "uva_head_prenet_forward", "uva_head_prenet_backward",
    "uva_head_sample", "uva_head_sample_from_decoder", "uva_decoder_tail_rows", "uva_head_loss_fwd", "uva_head_loss_bwd", "uva_head_launch_count",
    "uva_gemm_bf16_selftest", "uva_last_error", "uva_build_info", "uva_head_timing_enable", "uva_head_timing_read",
    "uva_adamw_ema_step", "uva_nccl_version", "uva_nccl_unique_id", "uva_nccl_comm_create", "uva_nccl_comm_destroy",
    "uva_head_allreduce_grads", "uva_head_allreduce_grads_bf16", "uva_head_set_grad_sync", "uva_head_grad_sync_launched", "uva_head_set_noise_seed", "uva_philox_normal", "uva_head_set_train_reserve",
]


def lib() -> ctypes.CDLL:
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if needs_build():
            build()
        L = ctypes.CDLL(os.environ.get("UVA_HEAD_LIB") or LIB_PATH)  # UVA_HEAD_LIB: an experimental build of the same sources
        vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_float
        L.uva_head_num_params.restype = i32
        L.uva_head_num_params.argtypes = [i32]
        L.uva_head_param_numel.restype = i64
        L.uva_head_param_numel.argtypes = [ctypes.POINTER(_Desc), i32]
        L.uva_head_create.restype = ctypes.c_int
        L.uva_head_create.argtypes = [ctypes.POINTER(_Desc), ctypes.POINTER(vp)]
        L.uva_head_destroy.restype = None
        L.uva_head_destroy.argtypes = [vp]
        L.uva_schedule_table.restype = ctypes.c_int
        L.uva_schedule_table.argtypes = [i32, i32, i32, ctypes.c_char_p, ctypes.POINTER(ctypes.c_double), i32]
        L.uva_head_load_weights.restype = ctypes.c_int
        L.uva_head_load_weights.argtypes = [vp, ctypes.POINTER(vp), i32, vp]
        L.uva_head_workspace_bytes.restype = ctypes.c_size_t
        L.uva_head_workspace_bytes.argtypes = [vp, i64, i32]
        L.uva_head_net_forward.restype = ctypes.c_int
        L.uva_head_net_forward.argtypes = [vp, vp, vp, vp, i64, vp, vp, ctypes.c_size_t, vp]
        L.uva_head_net_forward_static.restype = ctypes.c_int
        L.uva_head_net_forward_static.argtypes = [vp, vp, vp, vp, i64, vp, vp]
        L.uva_head_static_max_rows.restype = ctypes.c_int64
        L.uva_head_static_max_rows.argtypes = [vp]
        L.uva_head_sample.restype = ctypes.c_int
        L.uva_head_sample.argtypes = [vp, vp, vp, vp, vp, vp, f32, f32, i32, i32, i32, i64, vp, vp, ctypes.c_size_t, vp]
        L.uva_head_sample_from_decoder.restype = ctypes.c_int
        L.uva_head_sample_from_decoder.argtypes = [vp, vp, ctypes.POINTER(_Tail), vp, vp, vp, vp, vp, f32, f32, i32, i32, i32,
                                                   i64, vp, vp, ctypes.c_size_t, vp]
        L.uva_head_prenet_num_params.restype = i32
        L.uva_head_prenet_num_params.argtypes = []
        L.uva_head_prenet_load_weights.restype = ctypes.c_int
        L.uva_head_prenet_load_weights.argtypes = [vp, ctypes.POINTER(vp), i32, i32, i32, vp]
        L.uva_head_prenet_workspace_bytes.restype = ctypes.c_size_t
        L.uva_head_prenet_workspace_bytes.argtypes = [vp, i64, i32]
        L.uva_head_prenet_forward.restype = ctypes.c_int
        L.uva_head_prenet_forward.argtypes = [vp, vp, i64, vp, vp, ctypes.c_size_t, i32, vp]
        L.uva_head_prenet_backward.restype = ctypes.c_int
        L.uva_head_prenet_backward.argtypes = [vp, vp, i64, vp, ctypes.POINTER(vp), vp, ctypes.c_size_t, vp]
        L.uva_decoder_tail_rows.restype = ctypes.c_int
        L.uva_decoder_tail_rows.argtypes = [vp, ctypes.POINTER(_Tail), vp, i64, i32, vp, vp]
        L.uva_head_loss_fwd.restype = ctypes.c_int
        L.uva_head_loss_fwd.argtypes = [vp, vp, vp, vp, vp, vp, i64, vp, vp, vp, ctypes.c_size_t, vp]
        L.uva_head_loss_bwd.restype = ctypes.c_int
        L.uva_head_loss_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, i64, vp, ctypes.POINTER(vp), vp, ctypes.c_size_t, vp]
        L.uva_head_launch_count.restype = i64
        L.uva_head_launch_count.argtypes = [vp]
        L.uva_gemm_bf16_selftest.restype = ctypes.c_int
        L.uva_gemm_bf16_selftest.argtypes = [vp, vp, vp, vp, i64, i32, i32, i32, i32, vp]
        L.uva_head_timing_enable.restype = ctypes.c_int
        L.uva_head_timing_enable.argtypes = [vp, i32]
        L.uva_head_timing_read.restype = ctypes.c_int
        L.uva_head_timing_read.argtypes = [vp, i32, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(i64)]
        L.uva_adamw_ema_step.restype = ctypes.c_int
        L.uva_adamw_ema_step.argtypes = [i32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        L.uva_head_set_noise_seed.restype = ctypes.c_int
        L.uva_head_set_noise_seed.argtypes = [vp, ctypes.c_uint64, ctypes.c_uint32, i32]
        L.uva_philox_normal.restype = ctypes.c_int
        L.uva_philox_normal.argtypes = [ctypes.c_uint64, ctypes.c_uint32, i32, i64, i64, vp, vp]
        L.uva_head_set_train_reserve.restype = ctypes.c_int
        L.uva_head_set_train_reserve.argtypes = [vp, i32]
        L.uva_nccl_version.restype = i32
        L.uva_nccl_version.argtypes = []
        L.uva_nccl_unique_id.restype = ctypes.c_int
        L.uva_nccl_unique_id.argtypes = [vp]
        L.uva_nccl_comm_create.restype = ctypes.c_int
        L.uva_nccl_comm_create.argtypes = [i32, i32, vp, ctypes.POINTER(vp)]
        L.uva_nccl_comm_destroy.restype = ctypes.c_int
        L.uva_nccl_comm_destroy.argtypes = [vp]
        L.uva_head_allreduce_grads.restype = ctypes.c_int
        L.uva_head_allreduce_grads.argtypes = [vp, vp, i64, i32, vp]
        L.uva_head_set_grad_sync.restype = ctypes.c_int
        L.uva_head_set_grad_sync.argtypes = [vp, vp, vp, i32]
        L.uva_head_grad_sync_launched.restype = i32
        L.uva_head_grad_sync_launched.argtypes = [vp]
        L.uva_head_allreduce_grads_bf16.restype = ctypes.c_int
        L.uva_head_allreduce_grads_bf16.argtypes = [vp, vp, vp, i64, i32, vp]
        L.uva_last_error.restype = ctypes.c_char_p
        L.uva_build_info.restype = ctypes.c_char_p
        _lib = L
        return L


def _check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {lib().uva_last_error().decode()}")


def schedule_table(name: str, which: int, T_train: int = 1000, T_gen: int = 100):
    """float64 schedule table computed by the library's host code (no GPU needed)."""
    import numpy as np

    buf = (ctypes.c_double * T_train)()
    n = lib().uva_schedule_table(T_train, T_gen, which, name.encode(), buf, T_train)
    if n < 0:
        raise RuntimeError(lib().uva_last_error().decode())
    return np.array(buf[:n], dtype=np.float64)


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _f32c(t: torch.Tensor, what: str) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"{what}: expected a CUDA tensor (the B200 head has no CPU path)")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


class Head:
    """One diffusion head on the current CUDA device (opaque ``UvaHead*`` + a cached workspace)."""

    def __init__(self, C: int, Z: int, W: int, depth: int, T_train: int = 1000, T_gen: int = 100, mode: int = MODE_BF16):
        if not torch.cuda.is_available():
            raise RuntimeError("unified_video_action_b200 needs a CUDA (sm_100a) device; there is no CPU fallback")
        self.desc = _Desc(C, Z, W, depth, T_train, T_gen, mode)
        self.C, self.Z, self.W, self.depth, self.T_train, self.T_gen, self.mode = C, Z, W, depth, T_train, T_gen, mode
        h = ctypes.c_void_p()
        _check(lib().uva_head_create(ctypes.byref(self.desc), ctypes.byref(h)), "uva_head_create")
        self._h = h
        self.device = torch.device("cuda", torch.cuda.current_device())
        self._ws: Dict[int, torch.Tensor] = {}
        self._ws_streams: Dict[int, dict] = {}
        self.fwd_gen = 0  # bumped by every loss_fwd: whose activations the training workspace holds
        self.n_params = lib().uva_head_num_params(depth)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().uva_head_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # ---- plumbing -------------------------------------------------------------------------
    def workspace(self, rows: int, what: int) -> torch.Tensor:
        """Cached per-kind workspace.  A Head is single-stream and single-thread per call (like the reference
        modules), but successive calls may come from different streams (the action head's side stream, the eager
        backward stream) and the library's own second lane reads the buffer too: before a buffer is dropped for a
        larger one, every stream that used it is recorded on it so that the caching allocator does not hand the
        block out while kernels of those streams are still running."""
        need = int(lib().uva_head_workspace_bytes(self._h, rows, what))
        ws = self._ws.get(what)
        cur = torch.cuda.current_stream(self.device)
        users = self._ws_streams.setdefault(what, {})
        if ws is None or ws.numel() < need:
            if ws is not None:
                for st in users.values():
                    ws.record_stream(st)
                # the handle's internal lane stream is not a torch stream: drain it before the block can be reused
                torch.cuda.synchronize(self.device)
            users.clear()
            self._ws[what] = None
            ws = torch.empty(max(need, 1024), dtype=torch.uint8, device=self.device)
            self._ws[what] = ws
        users[cur.cuda_stream] = cur
        return ws

    def set_grad_sync(self, comm: Optional["GradComm"], average: bool = True) -> None:
        """While set, ``loss_bwd`` all-reduces ranges of the flat gradient buffer on the communicator's stream as it
        goes (``uva_head_set_grad_sync``); ``None`` turns it off."""
        if comm is None:
            _check(lib().uva_head_set_grad_sync(self._h, None, None, 1), "uva_head_set_grad_sync")
        else:
            _check(lib().uva_head_set_grad_sync(self._h, comm._c, comm.stream(self.device).cuda_stream, int(average)),
                   "uva_head_set_grad_sync")

    def grad_sync_launched(self) -> int:
        return int(lib().uva_head_grad_sync_launched(self._h))

    def set_train_reserve(self, sms: int) -> None:
        _check(lib().uva_head_set_train_reserve(self._h, int(sms)), "uva_head_set_train_reserve")

    def launch_count(self) -> int:
        return int(lib().uva_head_launch_count(self._h))

    def timing_enable(self, on: bool) -> None:
        _check(lib().uva_head_timing_enable(self._h, int(on)), "uva_head_timing_enable")

    def timing_read(self, category: int):
        """(total milliseconds, launches) of one GEMM category since the last read."""
        ms, n = ctypes.c_double(), ctypes.c_int64()
        _check(lib().uva_head_timing_read(self._h, category, ctypes.byref(ms), ctypes.byref(n)), "uva_head_timing_read")
        return ms.value, n.value

    def load_weights(self, params: Sequence[torch.Tensor]) -> None:
        if len(params) != self.n_params:
            raise RuntimeError(f"expected {self.n_params} parameter tensors, got {len(params)}")
        keep = [_f32c(p.detach(), "parameter") for p in params]
        for i, p in enumerate(keep):
            want = lib().uva_head_param_numel(ctypes.byref(self.desc), i)
            if p.numel() != want:
                raise RuntimeError(f"parameter {i}: {p.numel()} elements, expected {want}")
        arr = (ctypes.c_void_p * len(keep))(*[p.data_ptr() for p in keep])
        _check(lib().uva_head_load_weights(self._h, arr, len(keep), _stream()), "uva_head_load_weights")

    # ---- compute --------------------------------------------------------------------------
    def net_forward(self, x: torch.Tensor, t: torch.Tensor, c: torch.Tensor) -> torch.Tensor:
        x, c = _f32c(x, "x"), _f32c(c, "c")
        t = t.to(device=x.device, dtype=torch.int64).contiguous()
        N = x.shape[0]
        out = torch.empty(N, 2 * self.C, dtype=torch.float32, device=x.device)
        ws = self.workspace(N, 0)
        _check(lib().uva_head_net_forward(self._h, x.data_ptr(), t.data_ptr(), c.data_ptr(), N, out.data_ptr(),
                                          ws.data_ptr(), ws.numel(), _stream()), "uva_head_net_forward")
        return out

    def static_max_rows(self) -> int:
        return int(lib().uva_head_static_max_rows(self._h))

    def net_forward_static(self, x: torch.Tensor, t: torch.Tensor, c: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """``net_forward`` on the library's static device scratch (``uva_head_net_forward_static``): no workspace and no
        allocation inside the call - inputs must already be fp32 / int64, contiguous and on the device, ``out`` is the
        caller's ``[N, 2C]`` fp32 buffer.  At most ``static_max_rows()`` rows; concurrent calls race on the scratch."""
        for name, v, dt in (("x", x, torch.float32), ("c", c, torch.float32), ("t", t, torch.int64), ("out", out, torch.float32)):
            if v.dtype != dt or not v.is_contiguous() or not v.is_cuda:
                raise RuntimeError(f"net_forward_static: {name} must be a contiguous CUDA {dt} tensor")
        _check(lib().uva_head_net_forward_static(self._h, x.data_ptr(), t.data_ptr(), c.data_ptr(), x.shape[0],
                                                 out.data_ptr(), _stream()), "uva_head_net_forward_static")
        return out

    def sample(self, z, x_init, step_noise, temperature=1.0, cfg=1.0, clip_denoised=False, teacher=None,
               trace=False, k_begin=0, k_end=None, seed=None, offset=0):
        """``step_noise`` [steps, N, C], or ``None`` with ``seed`` (and a per-call ``offset``): the noise is
        then drawn in the kernel from Philox4x32-10 (no noise tensor in HBM)."""
        z, x_init = _f32c(z, "z"), _f32c(x_init, "x_init")
        N = z.shape[0]
        k_end = self.T_gen if k_end is None else k_end
        if step_noise is None:
            if seed is None:
                raise RuntimeError("sample: pass step_noise or a seed for the in-kernel Philox generator")
            _check(lib().uva_head_set_noise_seed(self._h, int(seed) & (2**64 - 1), int(offset) & 0xFFFFFFFF, 1),
                   "uva_head_set_noise_seed")
        else:
            step_noise = _f32c(step_noise, "step_noise")
            if step_noise.shape[0] < k_end - k_begin or step_noise.shape[1:] != x_init.shape:
                raise RuntimeError("step_noise must be [k_end-k_begin, N, C]")
        teacher = None if teacher is None else _f32c(teacher, "teacher")
        tr = torch.empty(k_end - k_begin, N, self.C, dtype=torch.float32, device=z.device) if trace else None
        out = torch.empty(N, self.C, dtype=torch.float32, device=z.device)
        if N == 0:
            return (out, tr) if trace else out
        ws = self.workspace(N, 0)
        _check(lib().uva_head_sample(self._h, z.data_ptr(), x_init.data_ptr(), _ptr(step_noise), _ptr(teacher),
                                     _ptr(tr), float(temperature), float(cfg), int(bool(clip_denoised)), int(k_begin),
                                     int(k_end), N, out.data_ptr(), ws.data_ptr(), ws.numel(), _stream()),
               "uva_head_sample")
        return (out, tr) if trace else out

    def sample_from_decoder(self, x_dec, tail: "DecoderTail", x_init, step_noise, temperature=1.0, cfg=1.0,
                            clip_denoised=False, row_index=None, trace=False, k_begin=0, k_end=None, seed=None, offset=0):
        """``sample`` with the condition rows taken from the MAR decoder output BEFORE ``decoder_norm``
        (``x_dec`` [B, L, Z]): LayerNorm + diffusion pos-embed run inside the library as the prologue of the
        cond_embed GEMM.  Rows: all B*T*S tokens, or ``row_index`` (int64 token-grid indices)."""
        x_dec, x_init = _f32c(x_dec, "x_dec"), _f32c(x_init, "x_init")
        if x_dec.dim() != 3 or x_dec.shape[2] != self.Z or tail.Z != self.Z:
            raise RuntimeError("x_dec must be [B, L, Z] with the head's Z")
        N = x_init.shape[0]
        k_end = self.T_gen if k_end is None else k_end
        if row_index is not None:
            row_index = row_index.to(device=x_dec.device, dtype=torch.int64).contiguous()
            if row_index.numel() != N:
                raise RuntimeError("row_index must have one entry per sampled row")
        if step_noise is None:
            if seed is None:
                raise RuntimeError("sample_from_decoder: pass step_noise or a seed for the in-kernel Philox generator")
            _check(lib().uva_head_set_noise_seed(self._h, int(seed) & (2**64 - 1), int(offset) & 0xFFFFFFFF, 1),
                   "uva_head_set_noise_seed")
        else:
            step_noise = _f32c(step_noise, "step_noise")
        tr = torch.empty(k_end - k_begin, N, self.C, dtype=torch.float32, device=x_dec.device) if trace else None
        out = torch.empty(N, self.C, dtype=torch.float32, device=x_dec.device)
        if N == 0:
            return (out, tr) if trace else out
        ws = self.workspace(N, 0)
        ts = tail.struct(x_dec.shape[1])
        _check(lib().uva_head_sample_from_decoder(self._h, x_dec.data_ptr(), ctypes.byref(ts), _ptr(row_index),
                                                  x_init.data_ptr(), _ptr(step_noise), None, _ptr(tr), float(temperature),
                                                  float(cfg), int(bool(clip_denoised)), int(k_begin), int(k_end), N,
                                                  out.data_ptr(), ws.data_ptr(), ws.numel(), _stream()),
               "uva_head_sample_from_decoder")
        return (out, tr) if trace else out

    # ---- fused DiffActLoss pre-net ----------------------------------------------------------
    def prenet_load_weights(self, params: Sequence[torch.Tensor], n_frames: int, n_actions: int) -> None:
        keep = [_f32c(p.detach(), "pre-net parameter") for p in params]
        arr = (ctypes.c_void_p * len(keep))(*[p.data_ptr() for p in keep])
        _check(lib().uva_head_prenet_load_weights(self._h, arr, len(keep), int(n_frames), int(n_actions), _stream()),
               "uva_head_prenet_load_weights")
        self.prenet_geom = (int(n_frames), int(n_actions))

    def prenet_workspace(self, B: int, train: bool) -> torch.Tensor:
        need = int(lib().uva_head_prenet_workspace_bytes(self._h, B, int(train)))
        key = 2
        ws = self._ws.get(key)
        cur = torch.cuda.current_stream(self.device)
        users = self._ws_streams.setdefault(key, {})
        if ws is None or ws.numel() < need:
            if ws is not None:
                for st in users.values():
                    ws.record_stream(st)
            users.clear()
            self._ws[key] = None
            ws = self._ws[key] = torch.empty(max(need, 1024), dtype=torch.uint8, device=self.device)
        users[cur.cuda_stream] = cur
        return ws

    def prenet_forward(self, z: torch.Tensor, train: bool = False) -> torch.Tensor:
        """z [B, n_frames * 256, Z] -> [B, n_actions, Z] (DAL:112-124)."""
        z = _f32c(z, "z")
        B = z.shape[0]
        T, J = self.prenet_geom
        if z.dim() != 3 or z.shape[1] != T * 256 or z.shape[2] != self.Z:
            raise RuntimeError(f"pre-net: z must be [B, {T * 256}, {self.Z}], got {tuple(z.shape)}")
        out = torch.empty(B, J, self.Z, dtype=torch.float32, device=z.device)
        ws = self.prenet_workspace(B, True) if train else self.prenet_workspace(B, False)
        # every forward rewrites the buffers a pending backward would read (an inference call shares them): the
        # generation counter lets that backward notice and redo its forward
        self.prenet_gen = getattr(self, "prenet_gen", 0) + 1
        _check(lib().uva_head_prenet_forward(self._h, z.data_ptr(), B, out.data_ptr(), ws.data_ptr(), ws.numel(), int(train),
                                             _stream()), "uva_head_prenet_forward")
        return out

    def prenet_backward(self, dout: torch.Tensor, B: int, dparams: Sequence[Optional[torch.Tensor]], want_dz: bool = True):
        dout = _f32c(dout, "dout")
        T, _ = self.prenet_geom
        dz = torch.empty(B, T * 256, self.Z, dtype=torch.float32, device=dout.device) if want_dz else None
        arr = (ctypes.c_void_p * len(dparams))(*[_ptr(g) for g in dparams])
        ws = self.prenet_workspace(B, True)
        _check(lib().uva_head_prenet_backward(self._h, dout.data_ptr(), B, _ptr(dz), arr, ws.data_ptr(), ws.numel(), _stream()),
               "uva_head_prenet_backward")
        return dz

    def loss_fwd(self, target, z, mask, t, noise, want_rows=False):
        target, z, noise = _f32c(target, "target"), _f32c(z, "z"), _f32c(noise, "noise")
        mask = None if mask is None else _f32c(mask, "mask")
        t = t.to(device=z.device, dtype=torch.int64).contiguous()
        N = z.shape[0]
        loss = torch.empty(1, dtype=torch.float32, device=z.device)
        rows = torch.empty(N, dtype=torch.float32, device=z.device) if want_rows else None
        ws = self.workspace(N, 1)
        self.fwd_gen += 1
        _check(lib().uva_head_loss_fwd(self._h, target.data_ptr(), z.data_ptr(), _ptr(mask), t.data_ptr(),
                                       noise.data_ptr(), N, loss.data_ptr(), _ptr(rows), ws.data_ptr(), ws.numel(),
                                       _stream()), "uva_head_loss_fwd")
        return (loss, rows) if want_rows else loss

    def loss_bwd(self, dloss, target, z, mask, t, noise, dparams: Sequence[Optional[torch.Tensor]], dz=None):
        target, z, noise = _f32c(target, "target"), _f32c(z, "z"), _f32c(noise, "noise")
        mask = None if mask is None else _f32c(mask, "mask")
        dloss = _f32c(dloss.reshape(1), "dloss")
        t = t.to(device=z.device, dtype=torch.int64).contiguous()
        N = z.shape[0]
        if dz is None:
            dz = torch.empty(N, self.Z, dtype=torch.float32, device=z.device)
        for g in dparams:
            if g is not None and (g.dtype != torch.float32 or not g.is_contiguous() or not g.is_cuda):
                raise RuntimeError("gradient buffers must be contiguous fp32 CUDA tensors")
        arr = (ctypes.c_void_p * len(dparams))(*[_ptr(g) for g in dparams])
        ws = self.workspace(N, 1)
        _check(lib().uva_head_loss_bwd(self._h, dloss.data_ptr(), target.data_ptr(), z.data_ptr(), _ptr(mask),
                                       t.data_ptr(), noise.data_ptr(), N, dz.data_ptr(), arr, ws.data_ptr(),
                                       ws.numel(), _stream()), "uva_head_loss_bwd")
        return dz


def decoder_tail_rows(x_dec: torch.Tensor, tail: DecoderTail, row_index: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 condition rows z [N, Z] = decoder_norm(x_dec)[:, skip:] + diffusion pos-embed (MAR:702-724) from the
    library's decoder-tail kernel (what ``sample_from_decoder`` feeds to cond_embed, before the bf16 rounding)."""
    x_dec = _f32c(x_dec, "x_dec")
    B, L, Z = x_dec.shape
    if row_index is not None:
        row_index = row_index.to(device=x_dec.device, dtype=torch.int64).contiguous()
    N = row_index.numel() if row_index is not None else B * tail.T * tail.S
    out = torch.empty(N, Z, dtype=torch.float32, device=x_dec.device)
    ts = tail.struct(L)
    with torch.cuda.device(x_dec.device):
        _check(lib().uva_decoder_tail_rows(x_dec.data_ptr(), ctypes.byref(ts), _ptr(row_index), N, Z, out.data_ptr(),
                                           _stream()), "uva_decoder_tail_rows")
    return out


def philox_normal(seed: int, offset: int, step0: int, per_step: int, steps: int, device) -> torch.Tensor:
    """[steps, per_step] standard normals of the sampler's in-kernel generator, written to memory."""
    out = torch.empty(steps, per_step, dtype=torch.float32, device=device)
    with torch.cuda.device(out.device):
        _check(lib().uva_philox_normal(int(seed) & (2**64 - 1), int(offset) & 0xFFFFFFFF, int(step0), int(per_step),
                                       out.numel(), out.data_ptr(), _stream()), "uva_philox_normal")
    return out


class GradComm:
    """NCCL communicator of the head-gradient all-reduce, created through the C ABI.  The
    128-byte unique id travels over the existing ``torch.distributed`` group (any backend); after that the
    collective itself is ``uva_head_allreduce_grads`` on the caller's CUDA stream."""

    def __init__(self, group=None):
        import torch.distributed as dist

        if not torch.cuda.is_available():
            raise RuntimeError("GradComm needs a CUDA device (NCCL); there is no CPU path in the library")
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        buf = ctypes.create_string_buffer(128)
        if self.rank == 0:
            _check(lib().uva_nccl_unique_id(buf), "uva_nccl_unique_id")
        box = [bytes(buf.raw)]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        c = ctypes.c_void_p()
        _check(lib().uva_nccl_comm_create(self.world, self.rank, ctypes.create_string_buffer(box[0], 128), ctypes.byref(c)),
               "uva_nccl_comm_create")
        self._c = c
        self.device = torch.device("cuda", torch.cuda.current_device())

    def stream(self, dev) -> "torch.cuda.Stream":
        """The one stream all collectives of this communicator run on (same order on every rank)."""
        if getattr(self, "_stream", None) is None:
            self._stream = torch.cuda.Stream(device=dev, priority=-1)
        return self._stream

    def all_reduce_async(self, flat: torch.Tensor, average: bool = True, bf16: bool = False) -> "torch.cuda.Event":
        """The same collective on this communicator's own stream, behind everything already enqueued on the
        caller's stream; returns the event a consumer of ``flat`` has to wait for.  All collectives of one
        communicator go through this one stream, so they keep one order on every rank."""
        dev = flat.device
        self.stream(dev)
        self._stream.wait_event(torch.cuda.current_stream(dev).record_event())
        with torch.cuda.stream(self._stream):
            self.all_reduce_(flat, average, bf16=bf16)
        flat.record_stream(self._stream)
        return self._stream.record_event()

    def all_reduce_(self, flat: torch.Tensor, average: bool = True, bf16: bool = False) -> torch.Tensor:
        """In-place all-reduce of a flat fp32 buffer on the current stream.  ``bf16``: the values cross NVLink as
        bf16 (half the bytes; rounded into a scratch buffer, reduced there, widened back)."""
        if not flat.is_cuda or flat.dtype != torch.float32 or not flat.is_contiguous():
            raise RuntimeError("GradComm.all_reduce_: expected a contiguous fp32 CUDA tensor")
        if bf16:
            scratch = torch.empty(flat.numel(), dtype=torch.bfloat16, device=flat.device)
            with torch.cuda.device(flat.device):
                _check(lib().uva_head_allreduce_grads_bf16(self._c, flat.data_ptr(), scratch.data_ptr(), flat.numel(),
                                                           int(average), _stream()), "uva_head_allreduce_grads_bf16")
            return flat
        with torch.cuda.device(flat.device):
            _check(lib().uva_head_allreduce_grads(self._c, flat.data_ptr(), flat.numel(), int(average), _stream()),
                   "uva_head_allreduce_grads")
        return flat

    def __del__(self):
        try:
            if getattr(self, "_c", None):
                lib().uva_nccl_comm_destroy(self._c)
                self._c = None
        except Exception:
            pass


def gemm_bf16_selftest(A: torch.Tensor, B: torch.Tensor, bias: Optional[torch.Tensor], block_n: int, layout: int = 0) -> torch.Tensor:
    """tcgen05 GEMM core unit check.  layout 0: A [M,K], B [N,K] -> A.B^T (+bias); 1: A [M,K], B [K,N] ->
    A.B; 2: A [K,M], B [K,N] -> A^T.B (split-K, reduced into a zero-filled D)."""
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16 and A.is_cuda and B.is_cuda
    A, B = A.contiguous(), B.contiguous()
    form = 0 if layout >= 24 else layout & 7  # 24, 25, 32, 33: forward form with a bf16 result (D holds bf16 [M, N])
    if form == 2:
        K, M = A.shape
    else:
        M, K = A.shape
    N = B.shape[0] if form == 0 else B.shape[1]
    D = torch.zeros(M, N, dtype=torch.float32, device=A.device)
    _check(lib().uva_gemm_bf16_selftest(A.data_ptr(), B.data_ptr(), _ptr(bias), D.data_ptr(), M, N, K, block_n, layout, _stream()),
           "uva_gemm_bf16_selftest")
    return D
