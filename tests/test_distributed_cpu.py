"""World-size-2 gloo tests (CPU) of the host-side multi-GPU logic:

* inference shards rows with no collective: the union of the ranks' shards is the whole batch, CFG
  shards keep cond / uncond rows paired;
* training all-reduces the head gradients once per head over a flat buffer and averages them.

The arithmetic of the head needs a B200; here the "gradients" are synthetic tensors on the reference's
parameter layout, which is all the collective code sees.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from unified_video_action_b200 import DiffActLoss, DiffLoss, allreduce_head_grads, shard_rows

        torch.manual_seed(0)  # same init on both ranks, like DDP's broadcast
        heads = [
            DiffLoss(16, 64, 2, 256, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None),
            DiffActLoss(2, 64, 2, 256, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None),
        ]
        # synthetic per-rank gradients: rank r holds (r + 1) * base
        g = torch.Generator().manual_seed(1)
        base = []
        for h in heads:
            for p in h.parameters():
                b = torch.randn(p.shape, generator=g)
                base.append(b)
                p.grad = (rank + 1.0) * b.clone()
        n = allreduce_head_grads(heads)
        assert n == len(heads)
        # the fused backward hands out views of one flat buffer: those are reduced in place (no packing)
        net_params = list(heads[0].net.parameters())
        flat = torch.zeros(sum((p.numel() + 3) // 4 * 4 for p in net_params))
        off, views = 0, []
        for p in net_params:
            v = flat[off:off + p.numel()].view(p.shape)
            v.copy_((rank + 1.0) * torch.ones_like(p))
            p.grad = v
            views.append(v)
            off += (p.numel() + 3) // 4 * 4
        assert allreduce_head_grads([heads[0]]) == 1
        for p, v in zip(net_params, views):
            assert p.grad.data_ptr() == v.data_ptr()
            torch.testing.assert_close(p.grad, torch.full_like(p, sum(r + 1.0 for r in range(world)) / world))
        # restore the synthetic gradients checked below
        i = 0
        for h in heads:
            for p in h.parameters():
                p.grad = (sum(r + 1.0 for r in range(world)) / world) * base[i].clone()
                i += 1
        want = sum(r + 1.0 for r in range(world)) / world
        i = 0
        for h in heads:
            for p in h.parameters():
                torch.testing.assert_close(p.grad, want * base[i], rtol=1e-6, atol=1e-6)
                i += 1

        # row sharding: plain and CFG-paired
        N = 1001
        b, e = shard_rows(N, rank, world)
        cover = torch.zeros(N)
        cover[b:e] = 1
        dist.all_reduce(cover)
        assert bool((cover == 1).all())
        N2 = 2 * 333
        b, e = shard_rows(N2, rank, world, pair_halves=True)
        cover = torch.zeros(N2)
        cover[b:e] = 1
        cover[N2 // 2 + b:N2 // 2 + e] = 1
        dist.all_reduce(cover)
        assert bool((cover == 1).all())
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_gloo_world2_allreduce_and_sharding(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert (tmp_path / f"ok{r}").exists()


class _FlatGrads(torch.autograd.Function):
    """Stand-in for ``_FusedLoss`` on CPU: hands out views of ONE flat buffer as the parameter gradients and arms
    the head's ``_GradSync`` exactly like ``_FusedLoss.backward`` does (the arithmetic of the real head needs a
    B200; the early all-reduce logic only sees the flat buffer and the autograd hooks)."""

    @staticmethod
    def forward(ctx, net, scale, *params):
        ctx.net, ctx.scale = net, scale
        return sum((p.detach() * p.detach()).sum() for p in params) * 0 + torch.tensor(1.0)

    @staticmethod
    def backward(ctx, dloss):
        from unified_video_action_b200.diffusion_loss import _grad_buffers

        params = list(ctx.net.parameters())
        flat, grads = _grad_buffers(params, torch.device("cpu"))
        for g, p in zip(grads, params):
            g.copy_(ctx.scale * torch.ones_like(p))
        sync = getattr(ctx.net, "_grad_sync", None)
        if sync is not None:
            sync.arm(flat, params)
        return (None, None, *grads)


def _accum_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from unified_video_action_b200 import allreduce_head_grads, overlap_allreduce

        class Head(torch.nn.Module):
            def __init__(self):
                super().__init__()
                self.net = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.Linear(7, 3))

        head = Head()
        overlap_allreduce([head])
        sync = head.net._grad_sync
        mean = sum(r + 1.0 for r in range(world)) / world

        def backward(scale):
            _FlatGrads.apply(head.net, scale, *head.net.parameters()).backward()

        # 1. one backward: autograd adopts the views, the hooks start the collective, the consumer only waits
        backward(rank + 1.0)
        assert sync.pending is not None and sync.armed is None
        assert allreduce_head_grads([head]) == 1
        for p in head.net.parameters():
            torch.testing.assert_close(p.grad, torch.full_like(p, mean))
        # 2. gradient accumulation: two backwards before ONE allreduce_head_grads.  The first
        # micro-step's early collective must be waited for and dropped, the accumulated sum reduced in full.
        for p in head.net.parameters():
            p.grad = None
        backward(rank + 1.0)
        assert sync.pending is not None
        backward(10.0 * (rank + 1.0))
        assert sync.pending is None, "the stale collective of micro-step 1 must not survive micro-step 2"
        assert allreduce_head_grads([head]) == 1
        for p in head.net.parameters():
            torch.testing.assert_close(p.grad, torch.full_like(p, 11.0 * mean))
        # 3. autograd clones instead of adopting (a tensor hook on one parameter): no early collective, plain reduce
        for p in head.net.parameters():
            p.grad = None
        hd = head.net[0].weight.register_hook(lambda g: g * 1.0)
        backward(rank + 1.0)
        hd.remove()
        assert sync.pending is None
        assert allreduce_head_grads([head]) >= 1
        for p in head.net.parameters():
            torch.testing.assert_close(p.grad, torch.full_like(p, mean))
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_gloo_world2_early_allreduce_under_gradient_accumulation(tmp_path):
    world = 2
    mp.spawn(_accum_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert (tmp_path / f"ok{r}").exists()
