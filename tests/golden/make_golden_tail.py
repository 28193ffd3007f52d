"""Golden vectors of the MAR decoder tail (mar_con_unified.py:702-724) -> tests/golden/decoder_tail.npz.

MAR itself cannot be imported in the build container (timm is missing), and the tail has no code of its own
beyond torch modules, so this script executes the reference's statements verbatim on the reference's module types:
`self.decoder_norm = norm_layer(decoder_embed_dim)` with `norm_layer=partial(nn.LayerNorm, eps=1e-6)` (MAR:252,
1201), the text-token strip (MAR:705-707) and the expand / add / reshape of the two diffusion embeddings
(MAR:709-722).  Run here once:  python tests/golden/make_golden_tail.py
"""
import os
from functools import partial

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    torch.manual_seed(0)
    B, T, S, D, buffer_size_text = 2, 4, 6, 768, 64
    norm_layer = partial(nn.LayerNorm, eps=1e-6)
    decoder_norm = norm_layer(D)
    with torch.no_grad():
        decoder_norm.weight.copy_(1.0 + 0.1 * torch.randn(D))
        decoder_norm.bias.copy_(0.1 * torch.randn(D))
    diffusion_temporal_embed = nn.Parameter(torch.zeros(1, T, D))
    diffusion_spatial_embed = nn.Parameter(torch.zeros(1, S, D))
    torch.nn.init.normal_(diffusion_temporal_embed, std=0.02)
    torch.nn.init.normal_(diffusion_spatial_embed, std=0.02)
    x_in = 3.0 * torch.randn(B, buffer_size_text + T * S, D) + 0.5
    embed_dim = D
    with torch.no_grad():
        # ---- statements of MAR.forward_mae_decoder, lines 702-724 ----
        x = decoder_norm(x_in)
        x = x[:, buffer_size_text:]
        diffusion_temporal_pos_embed_expanded = diffusion_temporal_embed.unsqueeze(2).expand(-1, -1, S, -1)
        diffusion_spatial_pos_embed_expanded = diffusion_spatial_embed.unsqueeze(1).expand(-1, T, -1, -1)
        diffusion_combined_pos_embed = (
            diffusion_temporal_pos_embed_expanded + diffusion_spatial_pos_embed_expanded
        ).reshape(1, T * S, embed_dim)
        x = x + diffusion_combined_pos_embed
    np.savez_compressed(
        os.path.join(HERE, "decoder_tail.npz"),
        x_dec=x_in.numpy(), ln_w=decoder_norm.weight.detach().numpy(), ln_b=decoder_norm.bias.detach().numpy(),
        temporal=diffusion_temporal_embed.detach().numpy(), spatial=diffusion_spatial_embed.detach().numpy(),
        z=x.numpy(), meta=np.array([B, T, S, D, buffer_size_text]), eps=np.array([1e-6]),
    )
    print("wrote decoder_tail.npz", x.shape)


if __name__ == "__main__":
    main()
