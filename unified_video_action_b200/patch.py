"""Glue for the callers of the head: in-place swap inside a reference ``MAR`` (MAR:269-344), row
sharding for multi-GPU sampling and the head-gradient all-reduce for data-parallel training
(SURVEY.md §8e)."""
from __future__ import annotations

from typing import Iterable, Optional, Tuple

import torch
import torch.nn as nn

from .diffusion_action_loss import DiffActLoss
from .diffusion_loss import DiffLoss

_HEAD_ATTRS = ("diffloss", "diffloss_wrist", "diffactloss", "diffproploss")


def _geometry(ref_head) -> dict:
    net = ref_head.net
    return dict(
        target_channels=ref_head.in_channels,
        z_channels=net.cond_embed.in_features,
        depth=len(net.res_blocks),
        width=net.cond_embed.out_features,
    )


def patch(mar_model: nn.Module, precision: Optional[str] = None) -> nn.Module:
    """Replace ``diffloss`` / ``diffloss_wrist`` / ``diffactloss`` / ``diffproploss`` of a reference
    MAR by the B200 heads, copying the weights (state_dict keys are identical)."""
    for attr in _HEAD_ATTRS:
        ref = getattr(mar_model, attr, None)
        if ref is None or isinstance(ref, (DiffLoss, DiffActLoss)):
            continue
        g = _geometry(ref)
        kw = dict(n_frames=getattr(ref, "n_frames", 4), language_emb_model=getattr(ref, "language_emb_model", None),
                  language_emb_model_type=getattr(ref, "language_emb_model_type", None), precision=precision)
        steps = str(ref.gen_diffusion.num_timesteps)
        if hasattr(ref, "act_model_type"):
            new = DiffActLoss(num_sampling_steps=steps, act_diff_training_steps=ref.train_diffusion.num_timesteps,
                              act_diff_testing_steps=steps, act_model_type=ref.act_model_type, **g, **kw)
        else:
            new = DiffLoss(num_sampling_steps=steps, **g, **kw)
        new.load_state_dict(ref.state_dict(), strict=True)
        p = next(ref.parameters())
        new.to(device=p.device)
        setattr(mar_model, attr, new)
    # the small heads run beside the video head on a side stream (MAR.forward_loss calls diffloss first
    # with the same z, MAR:770-784)
    vid = getattr(mar_model, "diffloss", None)
    if isinstance(vid, DiffLoss):
        for attr in ("diffactloss", "diffproploss"):
            small = getattr(mar_model, attr, None)
            if isinstance(small, DiffActLoss):
                small.overlap_with(vid)
    return mar_model


def shard_rows(n_rows: int, rank: int, world_size: int, pair_halves: bool = False) -> Tuple[int, int]:
    """[begin, end) of the rows rank ``rank`` owns.  Rows of the head are independent (DL:47-48), so
    inference shards by row with no collective.  With ``pair_halves`` (CFG: rows are cond || uncond)
    the split is over the first half; the caller takes the same range from both halves."""
    span = n_rows // 2 if pair_halves else n_rows
    base, rem = divmod(span, world_size)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


@torch.no_grad()
def sample_masked_tokens(head: DiffLoss, z: torch.Tensor, mask_to_pred: torch.Tensor, cur_tokens: torch.Tensor,
                         temperature: float = 1.0, cfg: float = 1.0, all_masked: Optional[bool] = None) -> torch.Tensor:
    """The glue around ``diffloss.sample`` inside ``MAR.sample_tokens`` (mar_con_unified.py:1086-1116) without its
    host round trips.  The reference gathers the rows to predict with ``z[mask_to_pred.nonzero(as_tuple=True)]``
    (a device-to-host sync for the row count), samples them, and scatters the result back with a second
    ``nonzero``.  Here: z [B, S, Z], mask_to_pred [B, S] (bool), cur_tokens [B, S, C] -> new tokens [B, S, C].

    ``all_masked=True`` (what every shipped config does: ``num_iter = 1`` predicts all tokens in one pass, uva.yaml:42)
    samples all B*S rows - the same rows in the same order as the reference's gather, hence the same RNG draws and
    the same samples - and writes them with one ``torch.where``: no ``nonzero``, no sync.  ``all_masked=None`` checks
    the mask (one sync, still one fewer than the reference) and falls back to the reference's gather / scatter for a
    partial mask (``num_iter > 1``), so the RNG consumption stays the reference's in every case.  cfg != 1 follows the
    reference's row layout (cond || uncond halves of z; only the cond half is scattered, MAR:1105-1116)."""
    B, S = mask_to_pred.shape
    C = cur_tokens.shape[-1]
    if all_masked is None:
        all_masked = bool(mask_to_pred.all())
    flat = cur_tokens.reshape(B * S, C)
    if all_masked:
        rows = z.reshape(-1, z.shape[-1])
        sampled = head.sample(rows, temperature, cfg)
        if not cfg == 1.0:
            sampled = sampled[: B * S]
        return torch.where(mask_to_pred.reshape(B * S, 1), sampled.to(flat.dtype), flat).reshape(B, S, C)
    m = mask_to_pred if cfg == 1.0 else torch.cat([mask_to_pred, mask_to_pred], dim=0)
    rows = z[m.nonzero(as_tuple=True)]
    sampled = head.sample(rows, temperature, cfg)
    if not cfg == 1.0:
        sampled, _ = sampled.chunk(2, dim=0)
    out = flat.clone().reshape(B, S, C)
    out[mask_to_pred.nonzero(as_tuple=True)] = sampled.to(out.dtype)
    return out


def _shared_flat(grads):
    """If every gradient is a contiguous view into one storage (what ``_FusedLoss.backward`` produces),
    return a 1-D tensor over the span they occupy; else None.  Gaps of the span hold alignment padding
    (zeros), which an all-reduce leaves harmlessly summed."""
    g0 = grads[0]
    st = g0.untyped_storage()
    lo, hi = None, None
    for g in grads:
        if g.dtype != g0.dtype or not g.is_contiguous() or g.untyped_storage().data_ptr() != st.data_ptr():
            return None
        b, e = g.storage_offset(), g.storage_offset() + g.numel()
        lo = b if lo is None else min(lo, b)
        hi = e if hi is None else max(hi, e)
    if hi - lo > 2 * sum(g.numel() for g in grads):
        return None
    return torch.empty(0, dtype=g0.dtype, device=g0.device).set_(st, lo, (hi - lo,))


_COMMS: dict = {}  # id(process group) -> _lib.GradComm (NCCL communicator created through the C ABI)


def _reduce_flat(flat: torch.Tensor, group, average: bool, world: int, bf16: bool = False) -> None:
    """One in-place all-reduce of a flat fp32 buffer.  CUDA buffers go through the library
    (``uva_head_allreduce_grads`` = ncclAllReduce with ncclAvg on the caller's stream); host buffers exist
    only in the CPU tests of this host logic and use the torch.distributed group (gloo) directly."""
    import torch.distributed as dist

    if flat.is_cuda:
        from . import _lib

        key = (id(group), flat.device.index)
        comm = _COMMS.get(key)
        if comm is None:
            comm = _COMMS[key] = _lib.GradComm(group)
        if getattr(comm, "_stream", None) is not None:
            # this communicator already runs collectives on its own stream (overlap_allreduce): keep them all there
            torch.cuda.current_stream(flat.device).wait_event(comm.all_reduce_async(flat, average, bf16=bf16))
        else:
            comm.all_reduce_(flat, average, bf16=bf16)
        return
    dist.all_reduce(flat, group=group)
    if average:
        flat /= world


class _Done:
    """Completion of one early collective: a CUDA event of the communicator's stream (NCCL through the C ABI) or
    a torch.distributed work handle (gloo, the CPU tests of this host logic)."""

    def __init__(self, event=None, work=None):
        self.event, self.work = event, work

    def wait(self, device) -> None:
        if self.event is not None:
            torch.cuda.current_stream(device).wait_event(self.event)
        if self.work is not None:
            self.work.wait()


class _GradSync:
    """Per-head state of the early all-reduce (``overlap_allreduce``).

    ``_FusedLoss.backward`` only ARMS it with the flat buffer whose views it returns as gradients; the collective
    itself is launched from the parameters' post-accumulate-grad hooks, when the last of them has fired and every
    ``p.grad`` really is a view of that buffer (autograd adopted the views).  If autograd cloned instead (hooks,
    ``create_graph``, a layout mismatch) or accumulated into an existing ``.grad`` (gradient accumulation), nothing
    is launched and ``allreduce_head_grads`` does the plain reduce.  A collective of an EARLIER backward that
    nobody consumed yet (second micro-step of an accumulation) is waited for in ``arm`` - before autograd adds the
    new gradients into the buffer that collective is still rewriting - and forgotten, so that
    ``allreduce_head_grads`` reduces the accumulated sum in full."""

    def __init__(self, group, average: bool, bf16: bool = False, in_backward: bool = False):
        self.group, self.average, self.bf16 = group, average, bf16
        # the library reduces gradient ranges from inside loss_bwd (fp32 wire only): the collective hides behind the
        # rest of the backward; see _FusedLoss.backward for when it applies
        self.in_backward = in_backward and not bf16
        self.strict = False  # set when a collective was launched before autograd had adopted the gradient views
        self.pending = None   # (flat, _Done) of the collective in flight
        self.armed = None     # flat buffer of the backward whose AccumulateGrad nodes are running now
        self.count = self.need = 0
        self.params = ()
        self.handles = []

    def comm(self, device):
        from . import _lib

        key = (id(self.group), device.index)
        c = _COMMS.get(key)
        if c is None:
            c = _COMMS[key] = _lib.GradComm(self.group)
        return c

    def arm(self, flat: torch.Tensor, params) -> None:
        if self.pending is not None:
            self.pending[1].wait(flat.device)
            self.pending = None
        self.armed, self.count, self.params = flat, 0, params
        self.need = sum(1 for p in params if p.requires_grad)

    def launched_in_backward(self, flat: torch.Tensor) -> None:
        """loss_bwd enqueued the range collectives itself: remember the buffer and the point of the communicator's
        stream a consumer has to wait for."""
        cs = self.comm(flat.device).stream(flat.device)
        flat.record_stream(cs)
        self.pending = (flat, _Done(event=cs.record_event()))
        self.armed, self.strict = None, True

    def hook(self, _param) -> None:
        if self.armed is None:
            return
        self.count += 1
        if self.count < self.need:
            return
        flat, self.armed = self.armed, None
        sp = flat.untyped_storage().data_ptr()
        if all(p.grad is not None and p.grad.untyped_storage().data_ptr() == sp for p in self.params if p.requires_grad):
            self.launch(flat)

    def launch(self, flat: torch.Tensor) -> None:
        if flat.is_cuda:
            self.pending = (flat, _Done(event=self.comm(flat.device).all_reduce_async(flat, self.average, bf16=self.bf16)))
            return
        import torch.distributed as dist

        if self.average:
            flat /= dist.get_world_size(self.group)
        self.pending = (flat, _Done(work=dist.all_reduce(flat, group=self.group, async_op=True)))


def overlap_allreduce(heads: Iterable[nn.Module], group=None, average: bool = True, bf16: bool = False,
                      in_backward: bool = False) -> None:
    """Start each head's gradient all-reduce as soon as its gradients are in place ("issued from the
    extension when the last wgrad tile lands, overlapped with the backbone's backward"): the fused backward of a
    head ends long before the MAR decoder / encoder backward does, so the collective runs on the communicator's own
    stream beside it.  ``allreduce_head_grads`` afterwards only waits for those collectives (and reduces what else
    the head owns, e.g. the PyTorch pre-net).  ``bf16``: the flat buffer crosses NVLink as bf16 (71.5 MB per head
    instead of 143 MB; converted on both sides by the library).  ``in_backward`` (fp32 wire): large heads are reduced
    range by range from INSIDE their fused backward (final layer, every block, every adaLN Linear, the embedders:
    3 * depth + 3 collectives, each started as soon as its gradients are enqueued), so the collective also hides behind
    the head's own backward; falls back to the hook-based start when gradients accumulate into existing ``.grad``, a
    parameter has tensor hooks or the batch is small.  Off by default: measured on 2 B200s it buys nothing in the
    head-only step (6.09 vs 6.02 ms p50; 1 GPU 5.5 ms) - the head's persistent GEMM grids fill every SM, NCCL's CTAs
    cannot co-reside with them and the collective ends up in the gaps either way; with fewer NCCL channels
    (NCCL_MAX_NCHANNELS 8 / 4 / 2: 6.5 / 8.4 / 12.4 ms) it only gets longer.  It is meant for a full training step,
    where the backbone's backward follows the heads'.  Needs the process group to be initialised;
    ``None`` heads or ``heads=[]`` are fine.  Call once after ``patch()``; calling it again replaces the hooks."""
    for head in heads:
        net = getattr(head, "net", None)
        if net is None:
            continue
        old = getattr(net, "_grad_sync", None)
        if old is not None:
            for hd in old.handles:
                hd.remove()
        sync = net._grad_sync = _GradSync(group, average, bf16, in_backward)
        for p in net.parameters():
            sync.handles.append(p.register_post_accumulate_grad_hook(sync.hook))


def allreduce_head_grads(heads: Iterable[nn.Module], group=None, average: bool = True, bf16: bool = False) -> int:
    """Data-parallel training: all-reduce of the head gradients over flat fp32 buffers (NCCL over NVLink
    through the C ABI on GPUs, gloo in the CPU tests): one collective per flat gradient buffer of the fused
    backward plus one for any other gradients of the head.  ``bf16``: the big flat buffers cross the link as bf16
    (half the bytes; the loose remainder stays fp32).  Returns the number of collectives issued."""
    import torch.distributed as dist

    n = 0
    world = dist.get_world_size(group)
    for head in heads:
        grads = [p.grad for p in head.parameters() if p.grad is not None]
        if not grads:
            continue
        # gradients that are views of one flat buffer (what the fused backward hands out) are reduced in
        # place; the others (e.g. the PyTorch pre-net of DiffActLoss) are packed into one buffer
        groups = {}
        for g in grads:
            groups.setdefault(g.untyped_storage().data_ptr(), []).append(g)
        # a collective that the backward already started on this head's flat buffer: only wait for it
        sync = getattr(getattr(head, "net", None), "_grad_sync", None)
        if sync is not None:
            sync.armed = None
        if sync is not None and sync.pending is not None:
            flat0, done = sync.pending
            sync.pending = None
            done.wait(flat0.device)  # always: nothing below may touch the buffer while the collective rewrites it
            if groups.pop(flat0.untyped_storage().data_ptr(), None) is not None:
                n += 1  # .grad are views of the reduced buffer
            elif sync.strict:
                raise RuntimeError("the head's gradients were all-reduced from inside its backward, but autograd did not keep "
                                   "the library's gradient buffer as .grad (cloned gradients would be torn): call "
                                   "overlap_allreduce(heads, in_backward=False)")
            sync.strict = False
        loose = []
        for members in groups.values():
            flat = _shared_flat(members) if len(members) > 1 else None
            if flat is None:
                loose.extend(members)
                continue
            _reduce_flat(flat, group, average, world, bf16)
            n += 1
        if loose:
            flat = torch.cat([g.reshape(-1) for g in loose])
            _reduce_flat(flat, group, average, world)
            off = 0
            for g in loose:
                g.copy_(flat[off:off + g.numel()].view_as(g))
                off += g.numel()
            n += 1
    return n
