"""Two-GPU test of the head-gradient all-reduce through the C ABI (uva_nccl_* / uva_head_allreduce_grads,
SURVEY.md 8e): two ranks run the same fused loss backward on different row shards; after
allreduce_head_grads every gradient must equal the mean of the two ranks' gradients (computed with a
plain torch.distributed all-reduce of copies) and the one-rank gradient of the concatenated batch scaled
as the mean of the two shard losses.  Skipped on boxes with fewer than two GPUs."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from unified_video_action_b200 import DiffLoss, _lib, allreduce_head_grads

        assert _lib.lib().uva_nccl_version() >= 21000
        # the raw collective: known values, sum and average
        comm = _lib.GradComm()
        x = torch.full((1000003,), float(rank + 1), device=dev)
        comm.all_reduce_(x, average=False)
        assert float(x.min()) == float(x.max()) == sum(r + 1.0 for r in range(world))
        comm.all_reduce_(x, average=True)
        assert float(x.min()) == float(x.max()) == sum(r + 1.0 for r in range(world))

        torch.manual_seed(0)  # same initial weights on every rank, like DDP's broadcast
        head = DiffLoss(16, 768, 6, 1024, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None,
                        precision="fp32")
        g = torch.Generator().manual_seed(1)
        with torch.no_grad():
            for p in head.net.parameters():
                if float(p.abs().max()) == 0.0:
                    p.copy_(0.02 * torch.randn(p.shape, generator=g))
        head = head.to(dev)
        gen = torch.Generator().manual_seed(100 + rank)
        B, S = 2, 96
        target = torch.randn(B, S, 16, generator=gen).to(dev)
        z = torch.randn(B, S, 768, generator=gen).to(dev).requires_grad_(True)
        mask = torch.ones(B, S, device=dev)
        torch.manual_seed(7 + rank)
        head(target, z, mask).backward()
        mine = {k: p.grad.clone() for k, p in head.net.named_parameters()}
        want = {}
        for k, v in mine.items():
            w = v.clone()
            dist.all_reduce(w)
            want[k] = w / world
        assert allreduce_head_grads([head]) == 1  # one flat buffer, one collective
        torch.cuda.synchronize()
        for k, p in head.net.named_parameters():
            scale = float(want[k].abs().max())
            assert float((p.grad - want[k]).abs().max()) <= 1e-6 * max(scale, 1e-30), k
        # the same step with the all-reduce started from inside the backward (overlap_allreduce)
        from unified_video_action_b200 import overlap_allreduce

        overlap_allreduce([head])
        for p in head.net.parameters():
            p.grad = None
        z.grad = None
        torch.manual_seed(7 + rank)
        head(target, z, mask).backward()
        assert head.net._grad_sync.pending is not None
        assert allreduce_head_grads([head]) == 1
        torch.cuda.synchronize()
        for k, p in head.net.named_parameters():
            scale = float(want[k].abs().max())
            assert float((p.grad - want[k]).abs().max()) <= 1e-5 * max(scale, 1e-30), ("early", k)
        # with gradient accumulation into existing .grad the early start must stand down
        torch.manual_seed(7 + rank)
        head(target, z, mask).backward()
        assert head.net._grad_sync.pending is None
        # gradient accumulation with the early all-reduce on: two backwards, ONE allreduce_head_grads.
        # Micro-step 1 starts a collective; micro-step 2 must wait for it before autograd adds into the same
        # buffer, drop it, and the final reduce must cover the accumulated sum: mean over ranks of (g1 + g2).
        for p in head.net.parameters():
            p.grad = None
        z.grad = None
        torch.manual_seed(7 + rank)
        head(target, z, mask).backward()
        assert head.net._grad_sync.pending is not None
        torch.manual_seed(7 + rank)
        head(target, z, mask).backward()
        assert head.net._grad_sync.pending is None
        assert allreduce_head_grads([head]) == 1
        torch.cuda.synchronize()
        for k, p in head.net.named_parameters():
            scale = float(want[k].abs().max())
            assert float((p.grad - 2.0 * want[k]).abs().max()) <= 2e-5 * max(scale, 1e-30), ("accumulate", k)
        # two forwards through the head before one backward: each backward differentiates ITS forward
        for p in head.net.parameters():
            p.grad = None
        z.grad = None
        torch.manual_seed(7 + rank)
        l1 = head(target, z, mask)
        l2 = head(0.5 * target, z, mask)  # overwrites the cached activations of l1
        l1.backward()
        allreduce_head_grads([head])
        torch.cuda.synchronize()
        for k, p in head.net.named_parameters():
            scale = float(want[k].abs().max())
            assert float((p.grad - want[k]).abs().max()) <= 2e-5 * max(scale, 1e-30), ("two forwards", k)
        del l2
        # bf16 on the wire (71.5 MB per head instead of 143 MB): same mean within bf16 rounding
        overlap_allreduce([head], bf16=True)
        for p in head.net.parameters():
            p.grad = None
        z.grad = None
        torch.manual_seed(7 + rank)
        head(target, z, mask).backward()
        assert allreduce_head_grads([head], bf16=True) == 1
        torch.cuda.synchronize()
        for k, p in head.net.named_parameters():
            scale = float(want[k].abs().max())
            assert float((p.grad - want[k]).abs().max()) <= 1e-2 * max(scale, 1e-30), ("bf16 wire", k)
        # a batch above the small-batch limit: the library reduces the gradient ranges from INSIDE loss_bwd
        # (uva_head_set_grad_sync: 3 * depth + 3 collectives on the communicator's stream, behind the running backward)
        gen = torch.Generator().manual_seed(200 + rank)
        B2, S2 = 2, 640
        target2 = torch.randn(B2, S2, 16, generator=gen).to(dev)
        z2 = torch.randn(B2, S2, 768, generator=gen).to(dev).requires_grad_(True)
        mask2 = (torch.rand(B2, S2, generator=gen) > 0.2).float().to(dev)
        overlap_allreduce([head], in_backward=False)
        for p in head.net.parameters():
            p.grad = None
        torch.manual_seed(17 + rank)
        head(target2, z2, mask2).backward()
        allreduce_head_grads([head])
        torch.cuda.synchronize()
        want2 = {k: p.grad.clone() for k, p in head.net.named_parameters()}
        overlap_allreduce([head], in_backward=True)
        for rep in range(2):
            for p in head.net.parameters():
                p.grad = None
            torch.manual_seed(17 + rank)
            head(target2, z2, mask2).backward()
            sync = head.net._grad_sync
            assert sync.pending is not None and sync.strict, "the in-backward collectives did not start"
            assert allreduce_head_grads([head]) == 1
            torch.cuda.synchronize()
            for k, p in head.net.named_parameters():
                scale = float(want2[k].abs().max())
                assert float((p.grad - want2[k]).abs().max()) <= 2e-5 * max(scale, 1e-30), ("in-backward", rep, k)
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_nccl_world2_head_gradient_allreduce(tmp_path):
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp

    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert (tmp_path / f"ok{r}").exists()
