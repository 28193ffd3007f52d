"""Fused AdamW + EMA step.

CPU: the oracle restatement is pinned to torch.optim.AdamW itself (the optimizer class the reference
instantiates, POL:343-360) and to the EMA recurrence of ema_model.py:58-89.
GPU: uva_adamw_ema_step against the oracle on the same seeded tensors (fp32: <= 1e-6 relative)."""
import copy

import pytest
import torch
import torch.nn as nn

from oracle import uva_oracle as O


def _toy(seed=0):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(37, 64), nn.LayerNorm(64), nn.Linear(64, 19))


def _groups(model, wd):
    """add_weight_decay (POL:325-341)."""
    decay, no_decay = [], []
    for name, p in model.named_parameters():
        (no_decay if (p.dim() == 1 or name.endswith(".bias")) else decay).append(p)
    return [{"params": no_decay, "weight_decay": 0.0}, {"params": decay, "weight_decay": wd}]


def test_oracle_matches_torch_adamw_and_ema_recurrence():
    lr, betas, eps, wd = 3e-3, (0.9, 0.95), 1e-8, 0.05
    ref = _toy()
    mine = copy.deepcopy(ref)
    opt = torch.optim.AdamW(_groups(ref, wd), lr=lr, betas=betas, eps=eps)
    names = [n for n, _ in mine.named_parameters()]
    P = [p.data for _, p in mine.named_parameters()]
    M = [torch.zeros_like(p) for p in P]
    V = [torch.zeros_like(p) for p in P]
    E = [p.clone() for p in P]
    E_ref = [p.detach().clone() for p in ref.parameters()]
    WD = [0.0 if (p.dim() == 1 or n.endswith(".bias")) else wd for n, p in zip(names, P)]
    g = torch.Generator().manual_seed(5)
    for step in range(1, 6):
        grads = [torch.randn(p.shape, generator=g) for p in P]
        for p, gr in zip(ref.parameters(), grads):
            p.grad = gr.clone()
        opt.step()
        decay = O.ema_get_decay(step - 1)
        for e, p in zip(E_ref, ref.parameters()):  # EMAModel.step, ema_model.py:82-86
            e.mul_(decay)
            e.add_(p.data, alpha=1 - decay)
        O.adamw_ema_step(P, grads, M, V, E, WD, lr, betas, eps, step, decay)
        for a, b in zip(P, ref.parameters()):
            torch.testing.assert_close(a, b.data, rtol=1e-6, atol=1e-7)
        for a, b in zip(E, E_ref):
            torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-7)
    # known answers of the warm-up schedule (ema_model.py:46-56): 0 for the first two steps, capped at max_value
    assert O.ema_get_decay(0) == 0.0 and O.ema_get_decay(1) == 0.0
    assert abs(O.ema_get_decay(2) - (1 - 2 ** (-2 / 3))) < 1e-12
    assert O.ema_get_decay(10 ** 9) == 0.9999
    from unified_video_action_b200 import ema_decay_at

    for s in (0, 1, 2, 17, 31623, 10 ** 7):
        assert ema_decay_at(s) == O.ema_get_decay(s)


@pytest.mark.gpu
def test_fused_step_matches_oracle_on_the_head():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    dev = torch.device("cuda", 0)
    from unified_video_action_b200 import DiffLoss, HeadAdamW

    torch.manual_seed(0)
    head = DiffLoss(16, 64, 2, 256, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None)
    with torch.no_grad():
        for p in head.parameters():
            p.copy_(0.05 * torch.randn(p.shape))
    ema_head = copy.deepcopy(head)
    names = [n for n, _ in head.named_parameters()]
    P = [p.detach().clone() for p in head.parameters()]
    E = [p.clone() for p in P]
    M = [torch.zeros_like(p) for p in P]
    V = [torch.zeros_like(p) for p in P]
    lr, betas, eps, wd = 1e-3, (0.9, 0.95), 1e-8, 0.02
    WD = [0.0 if (p.dim() == 1 or n.endswith(".bias")) else wd for n, p in zip(names, P)]
    head, ema_head = head.to(dev), ema_head.to(dev)
    opt = HeadAdamW([head], lr=lr, betas=betas, eps=eps, weight_decay=wd, ema_modules=[ema_head])
    g = torch.Generator().manual_seed(9)
    for step in range(1, 5):
        grads = [torch.randn(p.shape, generator=g) for p in P]
        for p, gr in zip(head.parameters(), grads):
            p.grad = gr.to(dev)
        opt.step(grad_scale=0.5)
        O.adamw_ema_step(P, grads, M, V, E, WD, lr, betas, eps, step, O.ema_get_decay(step - 1), grad_scale=0.5)
    for n, a, b in zip(names, head.parameters(), P):
        torch.testing.assert_close(a.detach().cpu(), b, rtol=2e-6, atol=1e-7, msg=n)
    for n, a, b in zip(names, ema_head.parameters(), E):
        torch.testing.assert_close(a.detach().cpu(), b, rtol=2e-6, atol=1e-7, msg=n)
    # the heads re-pack their operands after the step (parameters changed behind the version counters)
    assert head.net._packed_key is None
    z = torch.randn(300, 64, device=dev)
    out = head.sample(z, temperature=1.0, cfg=1.0)
    assert torch.isfinite(out).all()
    cpu_opt = HeadAdamW([copy.deepcopy(head).cpu()], lr=lr)  # construction is fine on CPU ...
    for p in cpu_opt.params:
        p.grad = torch.zeros_like(p)
    with pytest.raises(RuntimeError):
        cpu_opt.step()  # ... but stepping is CUDA only: no CPU fallback
