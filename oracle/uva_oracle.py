"""CPU oracle: a restatement of the reference's per-token diffusion head.  TEST INFRASTRUCTURE.

This file restates, op for op, the arithmetic of the reference hot path so that the CUDA
product path (``unified_video_action_b200``) can be checked against it.  It is *not* part of
the product: only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs
may import it.

Parity status: the reference ships no tests / golden vectors for this path, so
the oracle is pinned against the reference code itself: ``tests/golden/make_golden.py`` imports
the unmodified reference from ``/root/reference`` in the build container, runs it on seeded
inputs and commits the outputs under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks
this restatement against those vectors and against the float64 known-answer constants of
SURVEY.md §8(c).

Reference files restated (paths relative to
``/root/reference/unified_video_action/model/autoregressive/``):

  DL  = diffusion_loss.py            GD = diffusion/gaussian_diffusion.py
  DAL = diffusion_action_loss.py     RS = diffusion/respace.py
  DI  = diffusion/__init__.py        DU = diffusion/diffusion_utils.py

Arithmetic runs on torch CPU tensors (the reference is torch eager, so the same ATen ops in the
same order give bit-identical fp32 results); schedule tables are numpy float64 exactly like the
reference.  ``dtype=torch.float64`` gives a high-precision ground truth, and
``emulate_bf16=True`` rounds GEMM operands to bf16 at the points where the B200 bf16 kernels do,
so the bf16 kernels can be checked to ~1e-5 instead of only to the loose 1e-2 product tolerance.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------------------
# Schedules (numpy float64)                                           GD:102-202, RS:12-91
# --------------------------------------------------------------------------------------


def cosine_betas(num_steps: int, max_beta: float = 0.999) -> np.ndarray:
    """``get_named_beta_schedule("cosine")`` -> ``betas_for_alpha_bar``  (GD:102-145)."""

    def alpha_bar(t):
        return math.cos((t + 0.008) / 1.008 * math.pi / 2) ** 2

    betas = []
    for i in range(num_steps):
        t1 = i / num_steps
        t2 = (i + 1) / num_steps
        betas.append(min(1 - alpha_bar(t2) / alpha_bar(t1), max_beta))
    return np.array(betas, dtype=np.float64)


def space_timesteps(num_timesteps: int, section_counts) -> List[int]:
    """Sorted list of retained timesteps (RS:12-62).  'ddimN' strides are never used by the head."""
    if isinstance(section_counts, str):
        section_counts = [int(x) for x in section_counts.split(",")]
    size_per = num_timesteps // len(section_counts)
    extra = num_timesteps % len(section_counts)
    start_idx = 0
    all_steps = []
    for i, section_count in enumerate(section_counts):
        size = size_per + (1 if i < extra else 0)
        if size < section_count:
            raise ValueError(f"cannot divide section of {size} steps into {section_count}")
        frac_stride = 1 if section_count <= 1 else (size - 1) / (section_count - 1)
        cur_idx = 0.0
        for _ in range(section_count):
            all_steps.append(start_idx + round(cur_idx))
            cur_idx += frac_stride
        start_idx += size
    return sorted(set(all_steps))


@dataclass
class Schedule:
    """All float64 tables of ``GaussianDiffusion.__init__`` (GD:157-202) plus the respacing map."""

    betas: np.ndarray
    timestep_map: np.ndarray  # index in this (possibly respaced) process -> original timestep
    alphas_cumprod: np.ndarray
    alphas_cumprod_prev: np.ndarray
    sqrt_alphas_cumprod: np.ndarray
    sqrt_one_minus_alphas_cumprod: np.ndarray
    sqrt_recip_alphas_cumprod: np.ndarray
    sqrt_recipm1_alphas_cumprod: np.ndarray
    posterior_variance: np.ndarray
    posterior_log_variance_clipped: np.ndarray
    posterior_mean_coef1: np.ndarray
    posterior_mean_coef2: np.ndarray
    log_betas: np.ndarray

    @property
    def num_timesteps(self) -> int:
        return int(self.betas.shape[0])


def _tables_from_betas(betas: np.ndarray, timestep_map: Sequence[int]) -> Schedule:
    betas = np.array(betas, dtype=np.float64)
    assert betas.ndim == 1 and (betas > 0).all() and (betas <= 1).all()
    alphas = 1.0 - betas
    ac = np.cumprod(alphas, axis=0)
    ac_prev = np.append(1.0, ac[:-1])
    post_var = betas * (1.0 - ac_prev) / (1.0 - ac)
    post_logvar = np.log(np.append(post_var[1], post_var[1:])) if len(post_var) > 1 else np.array([])
    return Schedule(
        betas=betas,
        timestep_map=np.asarray(timestep_map, dtype=np.int64),
        alphas_cumprod=ac,
        alphas_cumprod_prev=ac_prev,
        sqrt_alphas_cumprod=np.sqrt(ac),
        sqrt_one_minus_alphas_cumprod=np.sqrt(1.0 - ac),
        sqrt_recip_alphas_cumprod=np.sqrt(1.0 / ac),
        sqrt_recipm1_alphas_cumprod=np.sqrt(1.0 / ac - 1),
        posterior_variance=post_var,
        posterior_log_variance_clipped=post_logvar,
        posterior_mean_coef1=betas * np.sqrt(ac_prev) / (1.0 - ac),
        posterior_mean_coef2=(1.0 - ac_prev) * np.sqrt(alphas) / (1.0 - ac),
        log_betas=np.log(betas),
    )


def create_schedule(timestep_respacing="", diffusion_steps: int = 1000) -> Schedule:
    """``create_diffusion(timestep_respacing, "cosine", diffusion_steps=...)`` (DI:11-47, RS:65-91)."""
    base_betas = cosine_betas(diffusion_steps)
    if timestep_respacing is None or timestep_respacing == "":
        timestep_respacing = [diffusion_steps]
    use = set(space_timesteps(diffusion_steps, timestep_respacing))
    base = _tables_from_betas(base_betas, list(range(diffusion_steps)))
    last = 1.0
    new_betas, tmap = [], []
    for i, a in enumerate(base.alphas_cumprod):
        if i in use:
            new_betas.append(1 - a / last)
            last = a
            tmap.append(i)
    return _tables_from_betas(np.array(new_betas), tmap)


def _extract(arr: np.ndarray, t: torch.Tensor, shape, dtype) -> torch.Tensor:
    """``_extract_into_tensor`` (GD:892-904): float64 gather, cast, broadcast via + zeros."""
    res = torch.from_numpy(arr).to(device=t.device)[t]  # like the reference: th.from_numpy(arr).to(device=timesteps.device)
    res = res.float() if dtype == torch.float32 else res.to(dtype)
    while res.dim() < len(shape):
        res = res[..., None]
    return res + torch.zeros(shape, dtype=res.dtype, device=t.device)


# --------------------------------------------------------------------------------------
# Parameters (reference state_dict names, DL:192-259)
# --------------------------------------------------------------------------------------


def param_specs(C: int, Z: int, W: int, depth: int, freq: int = 256) -> List[Tuple[str, Tuple[int, ...]]]:
    """(name, shape) of ``SimpleMLPAdaLN`` in the reference's state_dict order (DL:222-235)."""
    specs = [
        ("time_embed.mlp.0.weight", (W, freq)),
        ("time_embed.mlp.0.bias", (W,)),
        ("time_embed.mlp.2.weight", (W, W)),
        ("time_embed.mlp.2.bias", (W,)),
        ("cond_embed.weight", (W, Z)),
        ("cond_embed.bias", (W,)),
        ("input_proj.weight", (W, C)),
        ("input_proj.bias", (W,)),
    ]
    for i in range(depth):
        p = f"res_blocks.{i}."
        specs += [
            (p + "in_ln.weight", (W,)),
            (p + "in_ln.bias", (W,)),
            (p + "mlp.0.weight", (W, W)),
            (p + "mlp.0.bias", (W,)),
            (p + "mlp.2.weight", (W, W)),
            (p + "mlp.2.bias", (W,)),
            (p + "adaLN_modulation.1.weight", (3 * W, W)),
            (p + "adaLN_modulation.1.bias", (3 * W,)),
        ]
    specs += [
        ("final_layer.linear.weight", (2 * C, W)),
        ("final_layer.linear.bias", (2 * C,)),
        ("final_layer.adaLN_modulation.1.weight", (2 * W, W)),
        ("final_layer.adaLN_modulation.1.bias", (2 * W,)),
    ]
    return specs


def make_params(C: int, Z: int, W: int, depth: int, seed: int) -> Dict[str, np.ndarray]:
    """Deterministic, platform-independent non-degenerate parameters (numpy PCG64).

    The reference zero-initialises every AdaLN modulation layer and the final linear (DL:250-259),
    which makes a fresh head output exactly 0; parity on such weights is vacuous.  Here every
    matrix is drawn U(-a, a) with the Xavier bound a = sqrt(6/(fan_in+fan_out)) (what DL:238-244
    does for the non-zeroed layers), AdaLN / final / time-MLP matrices are N(0, 0.02), LN affine
    is 1 + N(0, 0.1) / N(0, 0.1) and biases are N(0, 0.02), so every term of the forward matters.
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    out: Dict[str, np.ndarray] = {}
    for name, shape in param_specs(C, Z, W, depth):
        if name.endswith("in_ln.weight"):
            a = 1.0 + 0.1 * rng.standard_normal(shape)
        elif name.endswith("in_ln.bias"):
            a = 0.1 * rng.standard_normal(shape)
        elif name.endswith(".bias"):
            a = 0.02 * rng.standard_normal(shape)
        elif "adaLN_modulation" in name or name.startswith("final_layer.linear") or name.startswith("time_embed"):
            a = 0.02 * rng.standard_normal(shape)
        else:
            bound = math.sqrt(6.0 / (shape[0] + shape[1]))
            a = rng.uniform(-bound, bound, size=shape)
        out[name] = a.astype(np.float32)
    return out


def to_torch_params(params: Dict[str, np.ndarray], dtype=torch.float32, requires_grad=False) -> Dict[str, torch.Tensor]:
    out = {}
    for k, v in params.items():
        t = torch.from_numpy(np.ascontiguousarray(v)).to(dtype) if isinstance(v, np.ndarray) else v.detach().to(dtype).cpu()
        t = t.clone()
        t.requires_grad_(requires_grad)
        out[k] = t
    return out


def depth_of(P: Dict[str, torch.Tensor]) -> int:
    d = 0
    while f"res_blocks.{d}.in_ln.weight" in P:
        d += 1
    return d


# --------------------------------------------------------------------------------------
# Denoiser network (DL:93-293)
# --------------------------------------------------------------------------------------


def _bf16r(x: torch.Tensor) -> torch.Tensor:
    """Round to bf16 and back (models a bf16 tensor-core operand / a bf16 activation in HBM)."""
    return x.to(torch.bfloat16).to(x.dtype)


def _linear(x, w, b, emulate_bf16=False, round_out=False):
    """nn.Linear (addmm).  With ``emulate_bf16`` the operands are bf16-rounded, accumulation is
    wide, bias is added in fp32 — what the tcgen05 bf16 kernels do."""
    if emulate_bf16:
        y = F.linear(_bf16r(x).double(), _bf16r(w).double()).to(x.dtype) + b
        return _bf16r(y) if round_out else y
    return F.linear(x, w, b)


def timestep_embedding(t: torch.Tensor, dim: int = 256, max_period: int = 10000, dtype=torch.float32) -> torch.Tensor:
    """Sinusoidal embedding cos||sin (DL:110-134).  The reference computes it in fp32."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(start=0, end=half, dtype=torch.float32) / half).to(device=t.device)
    if dtype == torch.float64:
        freqs = freqs.double()
        args = t[:, None].double() * freqs[None]
    else:
        args = t[:, None].float() * freqs[None]
    emb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    if dim % 2:
        emb = torch.cat([emb, torch.zeros_like(emb[:, :1])], dim=-1)
    return emb


def time_embed(P, t, dtype=torch.float32) -> torch.Tensor:
    """TimestepEmbedder.forward (DL:136-139): Linear(256,W) -> SiLU -> Linear(W,W)."""
    f = timestep_embedding(t, P["time_embed.mlp.0.weight"].shape[1], dtype=dtype)
    h = F.silu(F.linear(f, P["time_embed.mlp.0.weight"], P["time_embed.mlp.0.bias"]))
    return F.linear(h, P["time_embed.mlp.2.weight"], P["time_embed.mlp.2.bias"])


def net_forward(P, x, t, c, emulate_bf16: bool = False, taps: Optional[dict] = None) -> torch.Tensor:
    """SimpleMLPAdaLN.forward (DL:261-283) with ResBlock (DL:163-167) and FinalLayer (DL:185-189).

    x [N,C], t [N] (original-process timesteps, integer), c [N,Z] -> [N,2C].
    ``taps`` (optional dict) receives named intermediates for layer-by-layer debugging.
    """
    dt = x.dtype
    W = P["input_proj.weight"].shape[0]
    eb = emulate_bf16
    h = F.linear(x, P["input_proj.weight"], P["input_proj.bias"])  # tiny K: kept fp32 in every mode
    temb = time_embed(P, t, dtype=dt)
    cemb = _linear(c, P["cond_embed.weight"], P["cond_embed.bias"], eb)
    y = temb + cemb
    ys = F.silu(y)  # first op of every adaLN_modulation Sequential (DL:158-160, 181-183)
    if eb:
        ys = _bf16r(ys)
    if taps is not None:
        taps["cemb"], taps["temb"], taps["ysilu"], taps["x0"] = cemb, temb, ys, h
    for i in range(depth_of(P)):
        p = f"res_blocks.{i}."
        # the bf16 kernels store the whole modulation (shift | scale | gate) as bf16
        mod = _linear(ys, P[p + "adaLN_modulation.1.weight"], P[p + "adaLN_modulation.1.bias"], eb, round_out=eb)
        shift, scale, gate = mod.chunk(3, dim=-1)
        n = F.layer_norm(h, (W,), P[p + "in_ln.weight"], P[p + "in_ln.bias"], 1e-6)
        a = n * (1 + scale) + shift  # modulate (DL:93-94)
        if eb:
            a = _bf16r(a)
        u = F.silu(_linear(a, P[p + "mlp.0.weight"], P[p + "mlp.0.bias"], eb))
        if eb:
            u = _bf16r(u)
        v = _linear(u, P[p + "mlp.2.weight"], P[p + "mlp.2.bias"], eb)
        h = h + gate * v
        if taps is not None:
            taps[f"a{i}"], taps[f"u{i}"], taps[f"x{i + 1}"], taps[f"gate{i}"] = a, u, h, gate
    mod = _linear(ys, P["final_layer.adaLN_modulation.1.weight"], P["final_layer.adaLN_modulation.1.bias"], eb, round_out=eb)
    shift, scale = mod.chunk(2, dim=-1)
    n = F.layer_norm(h, (W,), None, None, 1e-6)
    a = n * (1 + scale) + shift
    if eb:
        a = _bf16r(a)
    out = _linear(a, P["final_layer.linear.weight"], P["final_layer.linear.bias"], eb)
    if taps is not None:
        taps["af"], taps["out"] = a, out
    return out


def net_forward_with_cfg(P, x, t, c, cfg_scale: float, emulate_bf16: bool = False) -> torch.Tensor:
    """SimpleMLPAdaLN.forward_with_cfg (DL:285-293): rows are cond || uncond."""
    C = P["input_proj.weight"].shape[1]
    half = x[: len(x) // 2]
    combined = torch.cat([half, half], dim=0)
    out = net_forward(P, combined, t, c, emulate_bf16)
    eps, rest = out[:, :C], out[:, C:]
    cond_eps, uncond_eps = torch.split(eps, len(eps) // 2, dim=0)
    half_eps = uncond_eps + cfg_scale * (cond_eps - uncond_eps)
    eps = torch.cat([half_eps, half_eps], dim=0)
    return torch.cat([eps, rest], dim=1)


# --------------------------------------------------------------------------------------
# Diffusion math (GD:220-444, 713-818; DU:10-73)
# --------------------------------------------------------------------------------------


def q_sample(S: Schedule, x_start, t, noise):
    """GD:220-236."""
    dt = x_start.dtype
    return (
        _extract(S.sqrt_alphas_cumprod, t, x_start.shape, dt) * x_start
        + _extract(S.sqrt_one_minus_alphas_cumprod, t, x_start.shape, dt) * noise
    )


def q_posterior_mean_variance(S: Schedule, x_start, x_t, t):
    """GD:238-258."""
    dt = x_t.dtype
    mean = (
        _extract(S.posterior_mean_coef1, t, x_t.shape, dt) * x_start
        + _extract(S.posterior_mean_coef2, t, x_t.shape, dt) * x_t
    )
    return mean, _extract(S.posterior_log_variance_clipped, t, x_t.shape, dt)


def p_mean_variance(S: Schedule, model_output, x, t, clip_denoised: bool):
    """GD:260-346 for EPSILON / LEARNED_RANGE, given the (already wrapped) model output [N,2C]."""
    dt = x.dtype
    C = x.shape[1]
    eps, var_values = torch.split(model_output, C, dim=1)
    min_log = _extract(S.posterior_log_variance_clipped, t, x.shape, dt)
    max_log = _extract(S.log_betas, t, x.shape, dt)
    frac = (var_values + 1) / 2
    log_variance = frac * max_log + (1 - frac) * min_log
    pred_xstart = (
        _extract(S.sqrt_recip_alphas_cumprod, t, x.shape, dt) * x
        - _extract(S.sqrt_recipm1_alphas_cumprod, t, x.shape, dt) * eps
    )  # GD:348-353
    if clip_denoised:
        pred_xstart = pred_xstart.clamp(-1, 1)
    mean, _ = q_posterior_mean_variance(S, pred_xstart, x, t)
    return mean, log_variance, pred_xstart


def p_sample(S: Schedule, model_output, x, t, noise, clip_denoised: bool, temperature: float):
    """GD:395-444 given the model output and the step noise (``randn_like(x)`` in the reference)."""
    mean, log_variance, pred_xstart = p_mean_variance(S, model_output, x, t, clip_denoised)
    nonzero_mask = (t != 0).to(x.dtype).view(-1, *([1] * (x.dim() - 1)))
    sample = mean + nonzero_mask * torch.exp(0.5 * log_variance) * noise * temperature
    return sample, pred_xstart, mean, log_variance


def sample_loop(
    P,
    S: Schedule,
    z: torch.Tensor,
    x_init: torch.Tensor,
    step_noise: torch.Tensor,
    temperature: float = 1.0,
    cfg: float = 1.0,
    clip_denoised: bool = False,
    emulate_bf16: bool = False,
    trace: bool = False,
    teacher: Optional[torch.Tensor] = None,
    num_steps: Optional[int] = None,
):
    """``p_sample_loop`` (GD:446-541) as driven by ``DiffLoss.sample`` / ``DiffActLoss.sample``
    (DL:68-90, DAL:207-227), without the hard-coded ``.cuda()`` calls.

    z [N,Z]; x_init [N,C] (for cfg != 1 the caller passes cat([noise, noise]) like DL:71-72);
    step_noise [T,N,C], entry k is the ``randn_like`` drawn at loop iteration k (timestep index
    T-1-k).  ``teacher`` [T,N,C] (optional) replaces the carried state before every step
    (teacher-forced per-step parity).  Returns final sample (and per-step states if ``trace``).
    """
    T = S.num_timesteps
    n_run = T if num_steps is None else num_steps
    img = x_init
    tmap = torch.from_numpy(S.timestep_map).to(device=x_init.device)
    states = []
    with torch.no_grad():
        for k in range(n_run):
            i = T - 1 - k
            if teacher is not None:
                img = teacher[k]
            t = torch.full((img.shape[0],), i, dtype=torch.long, device=img.device)
            new_ts = tmap[t]  # _WrappedModel (RS:122-127)
            if cfg != 1.0:
                out = net_forward_with_cfg(P, img, new_ts, z, cfg, emulate_bf16)
            else:
                out = net_forward(P, img, new_ts, z, emulate_bf16)
            img, _, _, _ = p_sample(S, out, img, t, step_noise[k], clip_denoised, temperature)
            if trace:
                states.append(img)
    if trace:
        return img, torch.stack(states)
    return img


def normal_kl(mean1, logvar1, mean2, logvar2):
    """DU:10-36."""
    return 0.5 * (-1.0 + logvar2 - logvar1 + torch.exp(logvar1 - logvar2) + ((mean1 - mean2) ** 2) * torch.exp(-logvar2))


def approx_standard_normal_cdf(x):
    """DU:39-44."""
    return 0.5 * (1.0 + torch.tanh(np.sqrt(2.0 / np.pi) * (x + 0.044715 * torch.pow(x, 3))))


def discretized_gaussian_log_likelihood(x, means, log_scales):
    """DU:47-73."""
    centered_x = x - means
    inv_stdv = torch.exp(-log_scales)
    plus_in = inv_stdv * (centered_x + 1.0 / 255.0)
    cdf_plus = approx_standard_normal_cdf(plus_in)
    min_in = inv_stdv * (centered_x - 1.0 / 255.0)
    cdf_min = approx_standard_normal_cdf(min_in)
    log_cdf_plus = torch.log(cdf_plus.clamp(min=1e-12))
    log_one_minus_cdf_min = torch.log((1.0 - cdf_min).clamp(min=1e-12))
    cdf_delta = cdf_plus - cdf_min
    return torch.where(
        x < -0.999,
        log_cdf_plus,
        torch.where(x > 0.999, log_one_minus_cdf_min, torch.log(cdf_delta.clamp(min=1e-12))),
    )


def vb_terms_bpd(S: Schedule, frozen_out, x_start, x_t, t):
    """GD:713-744 with clip_denoised=False (GD:789-796)."""
    true_mean, true_logvar = q_posterior_mean_variance(S, x_start, x_t, t)
    mean, log_variance, _ = p_mean_variance(S, frozen_out, x_t, t, clip_denoised=False)
    kl = normal_kl(true_mean, true_logvar, mean, log_variance)
    kl = kl.mean(dim=1) / np.log(2.0)
    nll = -discretized_gaussian_log_likelihood(x_start, means=mean, log_scales=0.5 * log_variance)
    nll = nll.mean(dim=1) / np.log(2.0)
    return torch.where(t == 0, nll, kl)


def training_losses(P, S: Schedule, x_start, t, noise, c, emulate_bf16: bool = False):
    """GD:746-818 for LossType.MSE + LEARNED_RANGE + EPSILON.  Returns dict(loss, mse, vb) [N]."""
    C = x_start.shape[1]
    x_t = q_sample(S, x_start, t, noise)
    new_ts = torch.from_numpy(S.timestep_map).to(device=t.device)[t]  # identity map for the training process (RS:122-127)
    model_output = net_forward(P, x_t, new_ts, c, emulate_bf16)
    eps, var_values = torch.split(model_output, C, dim=1)
    frozen_out = torch.cat([eps.detach(), var_values], dim=1)
    vb = vb_terms_bpd(S, frozen_out, x_start, x_t, t)
    mse = ((noise - eps) ** 2).mean(dim=1)
    return {"loss": mse + vb, "mse": mse, "vb": vb, "x_t": x_t, "model_output": model_output}


def diffloss_forward(P, S: Schedule, target, z, mask, t, noise, emulate_bf16: bool = False):
    """DiffLoss.forward (DL:44-66) with the RNG draws (t at DL:51, noise at GD:761) passed in.
    target [N,C], z [N,Z], mask [N] or None."""
    loss = training_losses(P, S, target, t, noise, z, emulate_bf16)["loss"]
    if mask is not None:
        loss = (loss * mask).sum() / mask.sum()
    return loss.mean()


def diffactloss_forward(P, S: Schedule, target, z, t, noise, emulate_bf16: bool = False):
    """DiffActLoss.forward after its pre-net (DAL:145-166): plain mean over rows."""
    return training_losses(P, S, target, t, noise, z, emulate_bf16)["loss"].mean()


def prenet_conv_fc(Pp, z, n_frames: int = 4, side: int = 16, emulate_bf16: bool = False):
    """The "conv_fc" pre-net of DiffActLoss (DAL:41-61; forward DAL:112-124 == 169-181).  ``Pp``: the 12 pre-net
    tensors under the reference's names (conv.0.*, fc.0.*, fc.2.*, interpolate.*, refine.0.*, refine.2.*).
    z [B, n_frames * side * side, Z] -> [B, 16, Z].  ``emulate_bf16``: bf16-rounded GEMM operands with wide
    accumulation at the points where the fused B200 pre-net rounds (conv / Linear inputs and weights)."""
    B, S, Z = z.shape
    r = (lambda t: _bf16r(t)) if emulate_bf16 else (lambda t: t)
    x = z.reshape(B * n_frames, S // n_frames, Z)            # "b (t s) c -> (b t) s c"
    x = x.reshape(B * n_frames, side, side, Z)               # "b (w h) c -> b w h c"
    x = x.permute(0, 3, 1, 2)                                # "b w h c -> b c w h"
    x = F.conv2d(r(x), r(Pp["conv.0.weight"]), Pp["conv.0.bias"], stride=1, padding=1)
    x = F.adaptive_avg_pool2d(F.relu(x), (4, 4))
    x = x.reshape(x.shape[0], -1)                            # "b c w h -> b (c w h)"
    x = F.relu(F.linear(r(x), r(Pp["fc.0.weight"]), Pp["fc.0.bias"]))
    x = F.linear(r(x), r(Pp["fc.2.weight"]), Pp["fc.2.bias"])
    x = x.reshape(B, n_frames, Z).permute(0, 2, 1)           # "(b t) c -> b t c", permute(0, 2, 1)
    x = F.linear(x, Pp["interpolate.weight"], Pp["interpolate.bias"]).permute(0, 2, 1)
    x = F.relu(F.linear(r(x), r(Pp["refine.0.weight"]), Pp["refine.0.bias"]))
    return F.linear(r(x), r(Pp["refine.2.weight"]), Pp["refine.2.bias"])


def decoder_tail(x_dec, ln_w, ln_b, eps, temporal, spatial, skip: int = 0):
    """Tail of ``MAR.forward_mae_decoder`` (mar_con_unified.py:702-724): ``decoder_norm`` (nn.LayerNorm, eps 1e-6 for
    mar_base, MAR:1201), drop the ``skip`` leading text tokens (MAR:705-707), add the diffusion position embedding
    temporal[1,T,1,D] + spatial[1,1,S,D] reshaped to [1, T*S, D] (MAR:709-722).  x_dec [B, L, D] -> z [B, T*S, D]."""
    D = x_dec.shape[-1]
    x = F.layer_norm(x_dec, (D,), ln_w, ln_b, eps)
    x = x[:, skip:]
    T, S = temporal.shape[-2], spatial.shape[-2]
    te = temporal.reshape(1, T, 1, D).expand(-1, -1, S, -1)
    se = spatial.reshape(1, 1, S, D).expand(-1, T, -1, -1)
    return x + (te + se).reshape(1, T * S, D)


# --------------------------------------------------------------------------------------
# Work counts — used by bench.py and DESIGN.md
# --------------------------------------------------------------------------------------


def sampling_flops_per_row_step(C: int, W: int = 1024, depth: int = 6) -> int:
    """Algorithmic FLOP per row-step with time_embed / cond_embed hoisted."""
    macs = C * W + depth * (3 * W * W + 2 * W * W) + 2 * W * W + W * 2 * C
    return 2 * macs


def training_flops_per_row(C: int, Z: int = 768, W: int = 1024, depth: int = 6, freq: int = 256) -> int:
    """fwd+bwd FLOP per row = 3 x 2 x MACs as executed by the reference."""
    macs = C * W + freq * W + W * W + Z * W + depth * (5 * W * W) + 2 * W * W + W * 2 * C
    return 6 * macs


# ------------------------------------------------------------------------------------------------
# Optimizer / EMA step for the head parameters
# ------------------------------------------------------------------------------------------------
def ema_get_decay(optimization_step: int, update_after_step=0, inv_gamma=1.0, power=2 / 3, min_value=0.0,
                  max_value=0.9999) -> float:
    """EMAModel.get_decay (model/autoregressive/ema_model.py:46-56)."""
    step = max(0, optimization_step - update_after_step - 1)
    value = 1 - (1 + step / inv_gamma) ** -power
    if step <= 0:
        return 0.0
    return max(min_value, min(value, max_value))


def adamw_ema_step(params, grads, exp_avg, exp_avg_sq, ema, weight_decay, lr, betas, eps, step, ema_decay,
                   grad_scale=1.0):
    """One torch.optim.AdamW update (amsgrad=False; the optimizer the reference builds in
    policy/unified_video_action_policy.py:343-360 with the decay groups of :325-341), written out
    tensor by tensor in the order of torch's single-tensor implementation, followed by
    EMAModel.step (ema_model.py:58-89): ema = ema * decay + p * (1 - decay).  All lists are torch CPU
    fp32 tensors updated IN PLACE; ``step`` is the 1-based count of this update; ``ema`` may be None."""
    b1, b2 = betas
    bc1 = 1 - b1 ** step
    bc2_sqrt = (1 - b2 ** step) ** 0.5
    for i, p in enumerate(params):
        g = grads[i] * grad_scale
        p.mul_(1 - lr * weight_decay[i])
        exp_avg[i].lerp_(g, 1 - b1)
        exp_avg_sq[i].mul_(b2).addcmul_(g, g, value=1 - b2)
        denom = (exp_avg_sq[i].sqrt() / bc2_sqrt).add_(eps)
        p.addcdiv_(exp_avg[i], denom, value=-(lr / bc1))
        if ema is not None and ema[i] is not None:
            ema[i].mul_(ema_decay)
            ema[i].add_(p, alpha=1 - ema_decay)


# ----------------------------------------------------------------------------------------------
# Counter-based sampler noise ("step_noise NULL -> in-kernel Philox(seed, offset)").
# Not a reference code path (the reference calls th.randn_like, GD:431): this restates the library's own
# generator so that the kernel can be checked bit-for-bit on the integer part.  Philox4x32-10 is the
# published algorithm (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3", SC'11;
# Random123 1.09), pinned below by the known-answer vectors of Random123's kat_vectors file.
# ----------------------------------------------------------------------------------------------
PHILOX_M0, PHILOX_M1 = 0xD2511F53, 0xCD9E8D57
PHILOX_W0, PHILOX_W1 = 0x9E3779B9, 0xBB67AE85


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10: uint32 arrays (broadcastable) -> four uint32 arrays."""
    c = [np.asarray(v, dtype=np.uint64) & 0xFFFFFFFF for v in (c0, c1, c2, c3)]
    k0 = np.uint64(k0 & 0xFFFFFFFF)
    k1 = np.uint64(k1 & 0xFFFFFFFF)
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(PHILOX_M0) * c[0]
        p1 = np.uint64(PHILOX_M1) * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & mask
        hi1, lo1 = p1 >> np.uint64(32), p1 & mask
        c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
        k0 = (k0 + np.uint64(PHILOX_W0)) & mask
        k1 = (k1 + np.uint64(PHILOX_W1)) & mask
    return [v.astype(np.uint32) for v in c]


def philox_normal(seed: int, offset: int, step0: int, per_step: int, steps: int) -> np.ndarray:
    """[steps, per_step] float32 standard normals: element e of loop iteration k uses counter
    (e >> 2 low, e >> 2 high, k, offset), key = seed; lanes (0,1) and (2,3) are Box-Muller pairs on 24-bit
    uniforms u1 = ((a >> 8) + 1) / 2^24, u2 = (b >> 8) / 2^24 (elementwise.cuh: philox_normal)."""
    e = np.arange(per_step, dtype=np.uint64)
    blk = e >> np.uint64(2)
    lane = (e & np.uint64(3)).astype(np.int64)
    out = np.empty((steps, per_step), dtype=np.float32)
    for j in range(steps):
        r = philox4x32_10(blk & np.uint64(0xFFFFFFFF), blk >> np.uint64(32), np.uint64(step0 + j), np.uint64(offset),
                          seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
        r = np.stack(r, axis=0)  # [4, per_step]
        a = np.take_along_axis(r, (lane & 2)[None, :], axis=0)[0]
        b = np.take_along_axis(r, ((lane & 2) + 1)[None, :], axis=0)[0]
        u1 = ((a >> np.uint32(8)).astype(np.float32) + np.float32(1.0)) * np.float32(1.0 / 16777216.0)
        u2 = (b >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)
        rad = np.sqrt(np.float32(-2.0) * np.log(u1.astype(np.float64))).astype(np.float32)
        ang = 2.0 * np.pi * u2.astype(np.float64)
        out[j] = rad * np.where(lane & 1, np.sin(ang), np.cos(ang)).astype(np.float32)
    return out
