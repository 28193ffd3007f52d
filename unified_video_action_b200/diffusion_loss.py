"""Drop-in ``DiffLoss`` backed by libuva_head.so (B200 / sm_100a).

Mirrors the reference ``DiffLoss`` (diffusion_loss.py, "DL"): same constructor arguments
(DL:11-42), the same ``state_dict`` keys and shapes under ``net.*`` (DL:222-235), the same
``forward(target, z, mask, ...)`` (DL:44-66) and ``sample(z, temperature, cfg, ...)`` (DL:68-90)
signatures and the same RNG call order (CPU ``randn`` for the start noise, one CUDA ``randn`` per
denoise step, ``randint`` then ``randn`` in training).  All arithmetic runs in the CUDA library;
there is no PyTorch fallback.
"""
from __future__ import annotations

import os
import weakref
from typing import List, Optional

import torch
import torch.nn as nn

from . import _lib


# id(DiffLoss module) -> (weak reference to z, z._version, event recorded when its forward() was entered);
# kept outside the modules so that copy.deepcopy (EMA) and state_dict never see CUDA events
_ENTRY: dict = {}
_EAGER_STREAMS: dict = {}  # device index -> stream of the eagerly enqueued backward (see _FusedLoss)


def _mode_from(precision: Optional[str]) -> int:
    p = (precision or os.environ.get("UVA_B200_PRECISION", "bf16")).lower()
    if p in ("bf16", "bfloat16"):
        return _lib.MODE_BF16
    if p in ("fp32", "float32"):
        return _lib.MODE_FP32
    raise ValueError(f"precision must be 'bf16' or 'fp32', got {precision!r}")


def _require_cuda(t: torch.Tensor, what: str) -> None:
    if t.device.type != "cuda":
        raise RuntimeError(f"{what}: the B200 diffusion head runs on CUDA (sm_100a) only; there is no CPU fallback")


class _TimestepEmbedder(nn.Module):
    """Parameter holder with the names of the reference TimestepEmbedder (DL:97-139)."""

    def __init__(self, hidden_size: int, frequency_embedding_size: int = 256):
        super().__init__()
        self.mlp = nn.Sequential(
            nn.Linear(frequency_embedding_size, hidden_size, bias=True),
            nn.SiLU(),
            nn.Linear(hidden_size, hidden_size, bias=True),
        )
        self.frequency_embedding_size = frequency_embedding_size


class _ResBlock(nn.Module):
    """Parameter holder with the names of the reference ResBlock (DL:142-167)."""

    def __init__(self, channels: int):
        super().__init__()
        self.channels = channels
        self.in_ln = nn.LayerNorm(channels, eps=1e-6)
        self.mlp = nn.Sequential(
            nn.Linear(channels, channels, bias=True), nn.SiLU(), nn.Linear(channels, channels, bias=True)
        )
        self.adaLN_modulation = nn.Sequential(nn.SiLU(), nn.Linear(channels, 3 * channels, bias=True))


class _FinalLayer(nn.Module):
    """Parameter holder with the names of the reference FinalLayer (DL:170-189)."""

    def __init__(self, model_channels: int, out_channels: int):
        super().__init__()
        self.norm_final = nn.LayerNorm(model_channels, elementwise_affine=False, eps=1e-6)
        self.linear = nn.Linear(model_channels, out_channels, bias=True)
        self.adaLN_modulation = nn.Sequential(nn.SiLU(), nn.Linear(model_channels, 2 * model_channels, bias=True))


def param_names(depth: int) -> List[str]:
    """Parameter names of ``SimpleMLPAdaLN`` in the order the C ABI expects (reference
    state_dict order, DL:222-235)."""
    names = [
        "time_embed.mlp.0.weight", "time_embed.mlp.0.bias", "time_embed.mlp.2.weight", "time_embed.mlp.2.bias",
        "cond_embed.weight", "cond_embed.bias", "input_proj.weight", "input_proj.bias",
    ]
    for i in range(depth):
        p = f"res_blocks.{i}."
        names += [p + s for s in (
            "in_ln.weight", "in_ln.bias", "mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias",
            "adaLN_modulation.1.weight", "adaLN_modulation.1.bias")]
    names += ["final_layer.linear.weight", "final_layer.linear.bias",
              "final_layer.adaLN_modulation.1.weight", "final_layer.adaLN_modulation.1.bias"]
    return names


class SimpleMLPAdaLN(nn.Module):
    """The AdaLN-MLP denoiser (DL:192-293).  Holds the fp32 ``nn.Parameter``s under the reference
    names (so checkpoints, EMA and optimizers work unchanged) and evaluates through the CUDA head."""

    def __init__(self, in_channels, model_channels, out_channels, z_channels, num_res_blocks,
                 grad_checkpointing=False, precision: Optional[str] = None, T_train: int = 1000, T_gen: int = 100):
        super().__init__()
        if out_channels != 2 * in_channels:
            raise ValueError("the B200 head implements learn_sigma heads only: out_channels must be 2*in_channels")
        self.in_channels = in_channels
        self.model_channels = model_channels
        self.out_channels = out_channels
        self.z_channels = z_channels
        self.num_res_blocks = num_res_blocks
        self.grad_checkpointing = grad_checkpointing  # accepted; the fused backward keeps what it needs
        self.time_embed = _TimestepEmbedder(model_channels)
        self.cond_embed = nn.Linear(z_channels, model_channels)
        self.input_proj = nn.Linear(in_channels, model_channels)
        self.res_blocks = nn.ModuleList([_ResBlock(model_channels) for _ in range(num_res_blocks)])
        self.final_layer = _FinalLayer(model_channels, out_channels)
        self.initialize_weights()
        self._mode = _mode_from(precision)
        self._T = (T_train, T_gen)
        self._head: Optional[_lib.Head] = None
        self._packed_key = None
        self._names = param_names(num_res_blocks)
        self.eager_backward = False  # see _FusedLoss
        self.train_reserve_sms = 0  # SMs this head leaves free during training for a head on another stream (overlap_with)
        self._reserve_set = None
        self.device_noise: Optional[bool] = None  # True: in-kernel Philox step noise (default: UVA_B200_NOISE env, "torch")

    def initialize_weights(self):
        """Same scheme as DL:237-259: Xavier-uniform Linears with zero bias, N(0, 0.02) timestep MLP,
        zero AdaLN modulation layers and zero output layer."""
        def _basic_init(m):
            if isinstance(m, nn.Linear):
                torch.nn.init.xavier_uniform_(m.weight)
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)
        self.apply(_basic_init)
        nn.init.normal_(self.time_embed.mlp[0].weight, std=0.02)
        nn.init.normal_(self.time_embed.mlp[2].weight, std=0.02)
        for block in self.res_blocks:
            nn.init.constant_(block.adaLN_modulation[-1].weight, 0)
            nn.init.constant_(block.adaLN_modulation[-1].bias, 0)
        nn.init.constant_(self.final_layer.adaLN_modulation[-1].weight, 0)
        nn.init.constant_(self.final_layer.adaLN_modulation[-1].bias, 0)
        nn.init.constant_(self.final_layer.linear.weight, 0)
        nn.init.constant_(self.final_layer.linear.bias, 0)

    # ---- library plumbing -------------------------------------------------------------------
    def ordered_params(self) -> List[torch.nn.Parameter]:
        """The parameters in C-ABI order.  Cached: walking named_parameters() costs ~0.3 ms of Python per call and
        the training step asks five times; the cache is dropped when a Parameter object is replaced (``_apply`` with
        swapped tensors, manual re-assignment) — in-place updates (optimizer, load_state_dict, .to()) keep identity."""
        cached = self.__dict__.get("_ordered")
        if cached is not None and cached[0] is self.time_embed.mlp[0].weight and cached[-1] is self.final_layer.adaLN_modulation[1].bias \
                and cached[4] is self.cond_embed.weight:
            return cached
        d = dict(self.named_parameters())
        cached = [d[n] for n in self._names]
        self.__dict__["_ordered"] = cached
        return cached

    def head(self) -> _lib.Head:
        """The CUDA head with up-to-date packed weights (re-packed when any parameter changed:
        optimizer step, load_state_dict, .to())."""
        params = self.ordered_params()
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("the B200 diffusion head runs on CUDA only; move the module to a cuda device")
        with torch.cuda.device(dev):
            if self._head is None or self._head.device != dev:
                self._head = _lib.Head(self.in_channels, self.z_channels, self.model_channels, self.num_res_blocks,
                                       self._T[0], self._T[1], self._mode)
                self._packed_key = None
                self._reserve_set = None
            if self._reserve_set != self.train_reserve_sms:
                self._head.set_train_reserve(self.train_reserve_sms)
                self._reserve_set = self.train_reserve_sms
            key = tuple((p.data_ptr(), p._version) for p in params)
            if key != self._packed_key:
                self._head.load_weights(params)
                self._packed_key = key
        return self._head

    def forward(self, x, t, c):
        """DL:261-283.  Inference-only entry (no autograd graph); training goes through
        ``DiffLoss.forward`` whose fused loss has its own backward."""
        _require_cuda(x, "SimpleMLPAdaLN.forward")
        with torch.cuda.device(x.device):
            return self.head().net_forward(x, t, c)

    def __deepcopy__(self, memo):
        # EMA copies (copy.deepcopy) must not share the CUDA handle
        import copy

        cls = self.__class__
        new = cls.__new__(cls)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            if k in ("_head", "_packed_key", "_grad_sync", "_ordered", "_reserve_set"):
                new.__dict__[k] = None
            else:
                new.__dict__[k] = copy.deepcopy(v, memo)
        return new


def _grad_buffers(params, device):
    """One zero-filled flat fp32 buffer + one view per parameter (a single memset instead of one fill per tensor)."""
    sizes = [(p.numel() + 3) // 4 * 4 if p.requires_grad else 0 for p in params]
    flat = torch.zeros(sum(sizes), dtype=torch.float32, device=device)
    grads, off = [], 0
    for p, n in zip(params, sizes):
        grads.append(flat[off:off + p.numel()].view(p.shape) if p.requires_grad else None)
        off += n
    return flat, grads


class _FusedLoss(torch.autograd.Function):
    """training_losses + masked mean as one op with a hand-written backward (GD:746-818, DL:64-66).

    ``net.eager_backward`` (set by ``DiffActLoss.overlap_with`` on the big head of a joint step): the op's output is
    the scalar loss, so its backward is linear in the incoming scalar gradient.  The whole backward is then enqueued
    right behind the forward with d loss = 1 and ``backward()`` only scales the result by the real d loss.  Why:
    autograd runs the head that was called LAST first, so in the joint step the host spends ~1.3 ms enqueueing the
    small action head's backward before it gets to the 16 384-row head, whose GPU stream idles meanwhile."""

    @staticmethod
    def forward(ctx, net: SimpleMLPAdaLN, target, z, mask, t, noise, *params):
        head = net.head()
        loss = head.loss_fwd(target, z, mask, t, noise)
        # the activations the backward needs live in the head's one cached training workspace: remember which
        # forward wrote them last (two forwards before a backward, or an eval forward in between)
        ctx.fwd_gen = head.fwd_gen
        ctx.net = net
        ctx.has_mask = mask is not None
        ctx.z_dtype = z.dtype
        ctx.eager = None
        if net.__dict__.pop("_eager_now", False):  # decided by the caller, where the grad mode is still visible
            # on its own stream: whatever the caller enqueues next on its stream (the other heads' joins, the loss
            # sum and ITS backward, which the small head's backward waits for) must not queue up behind 3 ms of GEMMs
            plist = net.ordered_params()
            dev = z.device
            cur = torch.cuda.current_stream(dev)
            flat, grads = _grad_buffers(plist, dev)
            one = torch.ones(1, dtype=torch.float32, device=dev)
            dz = torch.empty(z.shape[0], z.shape[1], dtype=torch.float32, device=dev)
            es = _EAGER_STREAMS.get(dev.index)
            if es is None:
                es = _EAGER_STREAMS[dev.index] = torch.cuda.Stream(device=dev)
            es.wait_event(cur.record_event())
            with torch.cuda.stream(es):
                head.loss_bwd(one, target, z, mask, t, noise, grads, dz=dz)
            for tns in (flat, dz, one, target, z, t, noise) + ((mask,) if mask is not None else ()):
                tns.record_stream(es)
            ctx.eager = (flat, grads, dz, es.record_event())
            ctx.save_for_backward()
        else:
            ctx.save_for_backward(target, z, mask if mask is not None else torch.empty(0, device=z.device), t, noise)
        return loss.reshape(())

    @staticmethod
    def backward(ctx, dloss):
        net = ctx.net
        params = net.ordered_params()
        if ctx.eager is not None:
            flat, grads, dz, done = ctx.eager
            ctx.eager = None  # a second backward through this node would scale twice
            torch.cuda.current_stream(flat.device).wait_event(done)
            g = dloss.float().reshape(1)
            flat.mul_(g)
            dz = dz.mul_(g)
        else:
            if not ctx.saved_tensors:
                raise RuntimeError("DiffLoss: backward through an eagerly differentiated loss can run only once")
            target, z, mask, t, noise = ctx.saved_tensors
            head = net.head()
            if head.fwd_gen != ctx.fwd_gen:
                # another forward went through this head since ours (e.g. (l1 + l2).backward(), an eval forward
                # between forward and backward, a re-packed head): the workspace no longer holds our activations.
                # Recompute them from the saved inputs - same draws, same arithmetic, so the same gradients as
                # the reference's autograd gives - instead of differentiating the wrong forward.
                head.loss_fwd(target, z, mask if ctx.has_mask else None, t, noise)
            flat, grads = _grad_buffers(params, z.device)
            sync = getattr(net, "_grad_sync", None)
            # data-parallel training: let the library all-reduce the gradient ranges while the backward is still running
            # (uva_head_set_grad_sync).  Only when autograd is sure to adopt the views of `flat` as .grad: nothing to
            # accumulate into, no tensor hooks, no double backward.
            inb = (sync is not None and sync.in_backward and not torch.is_grad_enabled()
                   and all(p.grad is None and not p._backward_hooks for p in params if p.requires_grad))
            if inb:
                if sync.pending is not None:  # an unconsumed collective of an earlier backward: not this path
                    inb = False
                else:
                    head.set_grad_sync(sync.comm(z.device), sync.average)
            try:
                dz = head.loss_bwd(dloss.float(), target, z, mask if ctx.has_mask else None, t, noise, grads)
            finally:
                if inb:
                    launched = head.grad_sync_launched()
                    head.set_grad_sync(None)
            if inb and launched > 0:
                sync.launched_in_backward(flat)
                return (None, None, dz.to(ctx.z_dtype) if ctx.needs_input_grad[2] else None, None, None, None, *grads)
        sync = getattr(net, "_grad_sync", None)
        if sync is not None:
            # data-parallel training (patch.overlap_allreduce): the head's gradients are complete here, long before
            # the backbone's backward is.  The collective starts from the parameters' post-accumulate-grad hooks once
            # autograd has adopted the views of `flat` as .grad (see patch._GradSync); a collective of an earlier,
            # unconsumed backward (gradient accumulation) is waited for here, before autograd adds into its buffer.
            sync.arm(flat, params)
        return (None, None, dz.to(ctx.z_dtype) if ctx.needs_input_grad[2] else None, None, None, None, *grads)


class DiffLoss(nn.Module):
    """Diffusion Loss (video-latent head) — drop-in for the reference class (DL:8-90)."""

    def __init__(self, target_channels, z_channels, depth, width, num_sampling_steps,
                 grad_checkpointing=False, **kwargs):
        super().__init__()
        self.n_frames = kwargs["n_frames"]
        self.language_emb_model = kwargs["language_emb_model"]
        self.language_emb_model_type = kwargs["language_emb_model_type"]
        self.in_channels = target_channels
        self.num_timesteps = 1000  # train_diffusion.num_timesteps (DI:11-20 default diffusion_steps)
        self.net = SimpleMLPAdaLN(
            in_channels=target_channels, model_channels=width, out_channels=target_channels * 2,
            z_channels=z_channels, num_res_blocks=depth, grad_checkpointing=grad_checkpointing,
            precision=kwargs.get("precision"), T_train=self.num_timesteps, T_gen=int(num_sampling_steps),
        )
        self.num_sampling_steps = int(num_sampling_steps)
        self._record_entry = False  # set by DiffActLoss.overlap_with(): remember where forward() was entered

    def forward(self, target, z, mask=None, conf_score=None, text_latents=None):
        _require_cuda(z, "DiffLoss.forward")
        if self._record_entry:
            # fork point for a head that runs beside this one on a side stream (see DiffActLoss.overlap_with)
            _ENTRY[id(self)] = (weakref.ref(z), z._version, torch.cuda.current_stream(z.device).record_event())
        bsz, seq_len, _ = target.shape
        target = target.reshape(bsz * seq_len, -1)
        z = z.reshape(bsz * seq_len, -1)
        mask = mask.reshape(bsz * seq_len)  # like DL:49 a mask is required
        # same draws, same order as the reference: t (DL:51), then noise = randn_like(x_start) (GD:761)
        t = torch.randint(0, self.num_timesteps, (target.shape[0],), device=target.device)
        noise = torch.randn_like(target, dtype=torch.float32)
        params = self.net.ordered_params()
        if self.net.eager_backward and torch.is_grad_enabled() and (z.requires_grad or any(p.requires_grad for p in params)):
            self.net.__dict__["_eager_now"] = True
        with torch.cuda.device(z.device):
            return _FusedLoss.apply(self.net, target.float(), z, mask.float(), t, noise, *params)

    @torch.no_grad()
    def sample(self, z, temperature=1.0, cfg=1.0, text_latents=None):
        return sample_rows(self.net, z, self.in_channels, self.num_sampling_steps, temperature, cfg, clip_denoised=False)


    @torch.no_grad()
    def sample_from_decoder(self, x_dec, decoder_norm, temporal_embed, spatial_embed, temperature=1.0, cfg=1.0,
                            skip_tokens: int = 0, row_index=None):
        """``sample(z)`` with ``z = decoder_norm(x_dec)[:, skip_tokens:] + diffusion pos-embed`` (the tail of
        ``MAR.forward_mae_decoder``, mar_con_unified.py:702-724) computed inside the head: pass the decoder blocks'
        output ``x_dec`` [B, L, Z], the ``decoder_norm`` LayerNorm module and the two embedding parameters
        ([1,T,Z], [1,S,Z]).  Same RNG call order and the same samples as ``sample`` on the reference's z; the fp32 z
        is never written.  ``row_index``: token-grid indices of the rows to sample
        (``mask_to_pred.nonzero()`` flattened, MAR:1090); default all B*T*S tokens."""
        _require_cuda(x_dec, "DiffLoss.sample_from_decoder")
        tail = _lib.DecoderTail(decoder_norm.weight, decoder_norm.bias, decoder_norm.eps, temporal_embed, spatial_embed,
                                skip=skip_tokens)
        N = int(row_index.numel()) if row_index is not None else x_dec.shape[0] * tail.T * tail.S
        return sample_rows(self.net, None, self.in_channels, self.num_sampling_steps, temperature, cfg, clip_denoised=False,
                           decoder=(x_dec, tail, row_index, N))


_NOISE_GRAPHS: dict = {}  # (device index, steps, N, C) -> (CUDAGraph, static [steps, N, C] buffer), latency regime only


def _step_noise(dev, steps: int, N: int, C: int) -> torch.Tensor:
    """The ``steps`` per-step draws ``randn(N, C)`` of the reference loop (GD:431), in the reference's order on
    torch's CUDA generator.  For small batches (predict_action: N = 16) the 100 launches cost more than the
    whole denoise chain's arithmetic, so they are captured once per shape into a CUDA graph and replayed:
    torch registers the generator with the graph, a replay consumes the same Philox offsets as the eager
    calls would, hence the same numbers (tests/test_gpu_infer.py).  ``UVA_NOISE_GRAPH=0`` disables it."""
    if N * C > 32768 or steps < 8 or os.environ.get("UVA_NOISE_GRAPH", "1") == "0":
        out = torch.empty(steps, N, C, dtype=torch.float32, device=dev)
        for k in range(steps):
            torch.randn(N, C, out=out[k])
        return out
    key = (dev.index, steps, N, C)
    entry = _NOISE_GRAPHS.get(key)
    if entry is None:
        if len(_NOISE_GRAPHS) >= 8:
            _NOISE_GRAPHS.pop(next(iter(_NOISE_GRAPHS)))
        buf = torch.empty(steps, N, C, dtype=torch.float32, device=dev)
        graph = torch.cuda.CUDAGraph()
        gen = torch.cuda.default_generators[dev.index]
        state = gen.get_state()
        with torch.cuda.graph(graph):
            for k in range(steps):
                torch.randn(N, C, out=buf[k])
        gen.set_state(state)  # capture itself must not consume the stream
        entry = _NOISE_GRAPHS[key] = (graph, buf)
    graph, buf = entry
    graph.replay()
    return buf


@torch.no_grad()
def sample_rows(net: SimpleMLPAdaLN, z, C: int, steps: int, temperature: float, cfg: float, clip_denoised: bool,
                decoder=None):
    """The sampling tail shared by DiffLoss.sample (DL:68-90) and DiffActLoss.sample (DAL:207-227):
    start noise from the CPU generator moved to the GPU, then the 100-step DDPM loop.  ``decoder`` =
    (x_dec, DecoderTail, row_index, N): the condition rows come from the MAR decoder output instead of ``z``."""
    if decoder is not None:
        x_dec, tail, row_index, N = decoder
        dev = x_dec.device
    else:
        _require_cuda(z, "sample")
        dev = z.device
        N = z.shape[0]
    if not cfg == 1.0:
        noise = torch.randn(N // 2, C).to(dev)
        noise = torch.cat([noise, noise], dim=0)
    else:
        noise = torch.randn(N, C).to(dev)
    if (getattr(net, "device_noise", None) if getattr(net, "device_noise", None) is not None
            else os.environ.get("UVA_B200_NOISE", "torch") == "philox"):
        # production mode: per-step noise drawn inside the DDPM-update kernel (Philox4x32-10), no
        # [steps, N, C] tensor; the seed comes from torch's CPU generator so manual_seed still reproduces it
        seed = int(torch.randint(0, 2**62, (1,)).item())
        with torch.cuda.device(dev):
            if decoder is not None:
                return net.head().sample_from_decoder(x_dec, tail, noise, None, temperature=temperature, cfg=cfg,
                                                      clip_denoised=clip_denoised, row_index=row_index, seed=seed)
            return net.head().sample(z, noise, None, temperature=temperature, cfg=cfg, clip_denoised=clip_denoised, seed=seed)
    # one CUDA randn per step, as th.randn_like(x) inside p_sample (GD:431)
    step_noise = _step_noise(dev, steps, N, C)
    with torch.cuda.device(dev):
        if decoder is not None:
            return net.head().sample_from_decoder(x_dec, tail, noise, step_noise, temperature=temperature, cfg=cfg,
                                                  clip_denoised=clip_denoised, row_index=row_index)
        return net.head().sample(z, noise, step_noise, temperature=temperature, cfg=cfg, clip_denoised=clip_denoised)
