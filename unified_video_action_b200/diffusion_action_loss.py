"""Drop-in ``DiffActLoss`` (action / proprioception head) backed by libuva_head.so.

Mirrors the reference class (diffusion_action_loss.py, "DAL"): constructor DAL:12-107,
``forward(target, z, task_mode, text_latents)`` DAL:109-166, ``sample(z, temperature, cfg,
text_latents)`` DAL:168-232.  The pre-net that pools ``z [B,1024,Z]`` to ``[B,16,Z]`` runs once
per call and stays PyTorch (cuDNN / cuBLAS) with the reference's parameter names; the denoiser,
the loss and the 100-step loop run in the CUDA library.
"""
from __future__ import annotations

import os
import weakref
from typing import Optional

import torch
import torch.nn as nn

from . import _lib
from .diffusion_loss import _ENTRY, SimpleMLPAdaLN, _FusedLoss, _grad_buffers, _require_cuda, sample_rows

_SIDE_STREAMS: dict = {}  # device index -> high-priority stream shared by the action heads of that device
_PRENET_GRAPHS: dict = {}  # id(module) -> (key, CUDAGraph, static input, static output); see DiffActLoss._prenet_replay
PRENET_GRAPH_MAX_BATCH = 8  # latency regime only: larger batches are bound by the convolution itself, not by launches


PRENET_PARAMS = ("conv.0.weight", "conv.0.bias", "fc.0.weight", "fc.0.bias", "fc.2.weight", "fc.2.bias",
                 "interpolate.weight", "interpolate.bias", "refine.0.weight", "refine.0.bias", "refine.2.weight", "refine.2.bias")


class _FusedPrenet(torch.autograd.Function):
    """The "conv_fc" pre-net (DAL:112-124) as one op of the CUDA library: implicit-GEMM convolution
    with the ReLU + 4 x 4 average pool in its epilogue, tcgen05 Linears; hand-written backward on the same cores."""

    @staticmethod
    def forward(ctx, mod, z, *params):
        head = mod._prenet_head()
        train = any(ctx.needs_input_grad)  # (grad mode itself is off inside Function.forward)
        out = head.prenet_forward(z, train=train)
        ctx.mod, ctx.B, ctx.z_dtype = mod, z.shape[0], z.dtype
        ctx.gen = getattr(head, "prenet_gen", 0)
        ctx.save_for_backward(z)
        return out

    @staticmethod
    def backward(ctx, dout):
        mod = ctx.mod
        (z,) = ctx.saved_tensors
        head = mod._prenet_head()
        if getattr(head, "prenet_gen", 0) != ctx.gen:  # another training forward overwrote the kept activations: redo ours
            head.prenet_forward(z, train=True)
        params = mod._prenet_params()
        flat, grads = _grad_buffers(params, dout.device)
        dz = head.prenet_backward(dout, ctx.B, grads, want_dz=ctx.needs_input_grad[1])
        return (None, dz.to(ctx.z_dtype) if dz is not None else None, *grads)


class DiffActLoss(nn.Module):
    """Diffusion Loss for action chunks."""

    def __init__(self, target_channels, z_channels, depth, width, num_sampling_steps, grad_checkpointing=False,
                 n_frames=4, act_diff_training_steps=1000, act_diff_testing_steps="100", act_model_type="conv_fc",
                 **kwargs):
        super().__init__()
        self.in_channels = target_channels
        self.n_frames = n_frames
        self.language_emb_model = kwargs["language_emb_model"]
        self.language_emb_model_type = kwargs["language_emb_model_type"]
        self.act_model_type = act_model_type
        if int(act_diff_training_steps) != 1000:
            # the sampling process is always respaced from 1000 steps (DAL:104-106)
            raise NotImplementedError("the B200 head supports act_diff_training_steps=1000 (all shipped configs)")

        if act_model_type == "conv_fc":  # DAL:35-61
            self.w = 16
            self.h = 16
            self.num_frames = 4
            self.num_actions = 16
            self.conv = nn.Sequential(
                nn.Conv2d(z_channels, z_channels, kernel_size=3, stride=1, padding=1),
                nn.ReLU(),
                nn.AdaptiveAvgPool2d((4, 4)),
            )
            self.fc = nn.Sequential(
                nn.Linear(z_channels * 4 * 4, z_channels), nn.ReLU(), nn.Linear(z_channels, z_channels)
            )
            self.interpolate = nn.Linear(self.num_frames, self.num_actions)
            self.refine = nn.Sequential(nn.Linear(z_channels, z_channels), nn.ReLU(), nn.Linear(z_channels, z_channels))
        elif act_model_type == "conv_ori":  # DAL:63-72
            self.w = 16
            self.h = 16
            self.conv_transpose3d = nn.ConvTranspose3d(z_channels, z_channels, kernel_size=(4, 1, 1), stride=(4, 1, 1))
            self.avg_pool = nn.AvgPool3d(kernel_size=(1, self.w, self.h))
        elif act_model_type == "conv2":  # DAL:74-79
            self.conv = nn.Sequential(
                nn.Conv1d(1024, 256, kernel_size=7, padding=3), nn.ReLU(), nn.Conv1d(256, 16, kernel_size=7, padding=3)
            )
        elif act_model_type == "fc2":  # DAL:81-86
            self.fc = nn.Sequential(nn.Linear(1024, 256), nn.ReLU(), nn.Linear(256, 16))
        else:
            raise NotImplementedError

        self.num_timesteps = int(act_diff_training_steps)
        self.num_sampling_steps = int(act_diff_testing_steps)
        object.__setattr__(self, "_fork_from", None)  # plain attribute: not a sub-module, not in state_dict
        self.net = SimpleMLPAdaLN(
            in_channels=target_channels, model_channels=width, out_channels=target_channels * 2,
            z_channels=z_channels, num_res_blocks=depth, grad_checkpointing=grad_checkpointing,
            precision=kwargs.get("precision"), T_train=self.num_timesteps, T_gen=self.num_sampling_steps,
        )

    # ---- fused pre-net ------------------------------------------------------------------------------------
    def _prenet_params(self):
        cached = self.__dict__.get("_pn_params")
        if cached is not None and cached[0] is self.conv[0].weight and cached[-1] is self.refine[2].bias:
            return cached
        d = dict(self.named_parameters())
        cached = self.__dict__["_pn_params"] = [d[n] for n in PRENET_PARAMS]
        return cached

    def _prenet_head(self) -> "_lib.Head":
        """The action head's library handle with up-to-date packed pre-net operands."""
        params = self._prenet_params()
        with torch.cuda.device(params[0].device):
            head = self.net.head()
            key = (id(head),) + tuple((p.data_ptr(), p._version) for p in params)
            if key != self.__dict__.get("_pn_key"):
                head.prenet_load_weights(params, self.num_frames, self.num_actions)
                self.__dict__["_pn_key"] = key
        return head

    def _fused_prenet_ok(self, z: torch.Tensor) -> bool:
        return (self.act_model_type == "conv_fc" and z.is_cuda and z.dim() == 3 and z.shape[1] == self.num_frames * self.w * self.h
                and self.net._mode == _lib.MODE_BF16 and z.shape[2] % 256 == 0 and self.n_frames == self.num_frames
                and os.environ.get("UVA_FUSED_PRENET", "1") != "0")

    def prenet(self, z: torch.Tensor) -> torch.Tensor:
        """z [B, 1024, Z] -> [B, 16, Z]  (DAL:112-143 == DAL:169-200).  bf16 mode on CUDA: the fused library pre-net
        (``UVA_FUSED_PRENET=0`` or the fp32 parity mode: the PyTorch modules below, the reference's own op sequence)."""
        if self._fused_prenet_ok(z):
            with torch.cuda.device(z.device):
                return _FusedPrenet.apply(self, z, *self._prenet_params())
        return self.prenet_torch(z)

    def prenet_torch(self, z: torch.Tensor) -> torch.Tensor:
        if self.act_model_type == "conv_fc":
            B, S, Cz = z.shape
            s = S // self.n_frames
            x = z.reshape(B * self.n_frames, s, Cz).reshape(B * self.n_frames, self.w, s // self.w, Cz)
            x = x.permute(0, 3, 1, 2)  # (b t) c w h
            x = self.conv(x)
            x = self.fc(x.reshape(x.shape[0], -1))
            x = x.reshape(B, self.n_frames, -1).permute(0, 2, 1)
            x = self.interpolate(x).permute(0, 2, 1)
            return self.refine(x)
        if self.act_model_type == "conv_ori":
            B, S, Cz = z.shape
            s = S // self.n_frames
            x = z.reshape(B, self.n_frames, self.w, s // self.w, Cz).permute(0, 4, 1, 2, 3)  # b c t w h
            x = self.avg_pool(self.conv_transpose3d(x))
            return x.permute(0, 2, 3, 4, 1).reshape(B, -1, Cz)
        if self.act_model_type == "conv2":
            return self.conv(z)
        if self.act_model_type == "fc2":
            return self.fc(z.transpose(1, 2)).transpose(1, 2)
        raise NotImplementedError

    def _prenet_replay(self, z: torch.Tensor) -> torch.Tensor:
        """Opt-in (``UVA_PRENET_GRAPH=1``): the pre-net of one small inference call as a CUDA graph replay.  At
        B = 1 ``prenet`` is ~25 cuDNN / cuBLAS / elementwise launches (0.29 ms) in front of a 5.5 ms sampling call.
        Measured on B200: the replay does not shorten the call (p50 6.05 -> 6.08 ms) - the 0.29 ms are GPU time of
        library kernels that are badly shaped for 4 images (the Linear(12288, 768) alone streams 37.7 MB of fp32
        weight through an SGEMM), not launch overhead - so it stays off by default; what would help is the fused
        pre-net kernel of SURVEY 8f-1.  The graph reads the parameters where they live, so in-place updates
        (optimizer steps, load_state_dict) are seen; anything that moves them (``.to()``, a dtype cast), another
        shape or another autocast state re-captures.  The returned tensor is the graph's static output: valid
        until the next call on this module (``sample`` consumes it at once on the same stream)."""
        if self._fused_prenet_ok(z):  # the library pre-net is a handful of launches already: nothing to replay
            return self.prenet(z)
        if (not z.is_cuda or z.shape[0] > PRENET_GRAPH_MAX_BATCH or torch.is_grad_enabled()
                or os.environ.get("UVA_PRENET_GRAPH", "0") == "0" or torch.cuda.is_current_stream_capturing()):
            return self.prenet(z)
        dev = z.device
        params = [p for n, p in self.named_parameters() if not n.startswith("net.")]
        try:
            cast = (torch.is_autocast_enabled("cuda"), torch.get_autocast_dtype("cuda"))
        except TypeError:  # older torch: no device argument
            cast = (torch.is_autocast_enabled(), torch.get_autocast_gpu_dtype())
        key = (dev.index, tuple(z.shape), z.dtype, cast, tuple(p.data_ptr() for p in params), tuple(p.dtype for p in params))
        ent = _PRENET_GRAPHS.get(id(self))
        if ent is None or ent[0] != key:
            with torch.cuda.device(dev):
                cur = torch.cuda.current_stream(dev)
                warm = torch.cuda.Stream(device=dev)
                static_in = torch.empty(z.shape, dtype=z.dtype, device=dev)
                warm.wait_stream(cur)
                with torch.cuda.stream(warm):
                    static_in.copy_(z)
                    for _ in range(3):  # library heuristics, workspaces and lazy module loads happen outside the capture
                        self.prenet(static_in)
                cur.wait_stream(warm)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    static_out = self.prenet(static_in)
            ent = _PRENET_GRAPHS[id(self)] = (key, graph, static_in, static_out)
            weakref.finalize(self, _PRENET_GRAPHS.pop, id(self), None)
        _, graph, static_in, static_out = ent
        static_in.copy_(z)
        graph.replay()
        return static_out

    # ---- side-stream execution -------------------------------------------------------------
    def overlap_with(self, other: Optional[nn.Module]) -> "DiffActLoss":
        """Run this head beside ``other`` (the video ``DiffLoss`` that is called just before it with the
        same ``z``, MAR:770-778): forward() then forks its side stream from the point where
        ``other.forward`` was entered instead of from its own entry, so the 256-row action head and its
        pre-net fill the gaps of the 16 384-row video head in the forward pass as well as in the backward
        pass.  Requires that ``target`` of this head was complete on the stream before ``other`` was
        called (true in MAR.forward_loss, where both are function arguments).  ``None`` undoes it."""
        object.__setattr__(self, "_fork_from", None if other is None else weakref.ref(other))
        if other is not None:
            other._record_entry = True
            # the big head can leave SMs to this head's kernels (UVA_TRAIN_RESERVE_SMS; measured: no gain, default 0)
            other.net.train_reserve_sms = int(os.environ.get("UVA_TRAIN_RESERVE_SMS", "0"))
            # ... and can differentiate eagerly (its backward enqueued with its forward on a third stream, see
            # _FusedLoss; UVA_EAGER_BACKWARD=1).  Both were measured on the C2 step and change nothing (6.4 ms either
            # way: the two heads' kernels add up rather than overlap), so both stay off by default.
            other.net.eager_backward = os.environ.get("UVA_EAGER_BACKWARD", "0") != "0"
            quiet = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
            if quiet is not None:  # parameters are created on the default stream and used on the side stream on purpose
                quiet(False)
        return self

    @staticmethod
    def _side(dev) -> torch.cuda.Stream:
        s = _SIDE_STREAMS.get(dev.index)
        if s is None:
            s = _SIDE_STREAMS[dev.index] = torch.cuda.Stream(device=dev, priority=-1)
        return s

    def forward(self, target, z, task_mode=None, text_latents=None):
        """DAL:109-166.  With autograd on (training) the pre-net, the fused loss and therefore (autograd
        runs a backward node on the stream of its forward) their backward run on a high-priority side
        stream: the head has 16 rows per sample, so alone it is a chain of ~150 launch-bound kernels; next
        to the video head it costs nothing.  ``UVA_ACT_OVERLAP=0`` keeps everything on the caller's stream."""
        _require_cuda(z, "DiffActLoss.forward")
        if not torch.is_grad_enabled() or os.environ.get("UVA_ACT_OVERLAP", "1") == "0":
            return self._forward(target, z)
        dev = z.device
        cur = torch.cuda.current_stream(dev)
        before = self.net._packed_key
        with torch.cuda.device(dev):
            self.net.head()  # a re-pack after an optimizer step stays on the caller's stream
        ev = None
        src = self._fork_from() if self._fork_from is not None else None
        entry = _ENTRY.get(id(src)) if src is not None else None
        if entry is not None and self.net._packed_key == before:
            zref, ver, e = entry
            if zref() is z and ver == z._version:  # the very tensor the video head was given, unmodified since
                ev = e
        if ev is None:
            ev = cur.record_event()
        side = self._side(dev)
        side.wait_event(ev)
        with torch.cuda.stream(side):
            loss = self._forward(target, z)
        z.record_stream(side)
        target.record_stream(side)
        cur.wait_stream(side)
        loss.record_stream(cur)
        return loss

    def _forward(self, target, z):
        bsz, seq_len, _ = target.shape
        z = self.prenet(z)
        target = target.reshape(bsz * seq_len, -1)
        z = z.reshape(bsz * seq_len, -1)
        t = torch.randint(0, self.num_timesteps, (target.shape[0],), device=target.device)  # DAL:148-153
        noise = torch.randn_like(target, dtype=torch.float32)  # GD:761
        with torch.cuda.device(z.device):
            # plain mean over rows (DAL:160-164): mask = None
            return _FusedLoss.apply(self.net, target.float(), z, None, t, noise, *self.net.ordered_params())

    @torch.no_grad()
    def sample(self, z, temperature=1.0, cfg=1.0, text_latents=None):
        z = self._prenet_replay(z)
        bsz, seq_len, _ = z.shape
        z = z.reshape(bsz * seq_len, -1)
        out = sample_rows(self.net, z, self.in_channels, self.num_sampling_steps, temperature, cfg, clip_denoised=True)
        return out.reshape(bsz, seq_len, -1)
