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
