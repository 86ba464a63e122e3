"""Golden vectors for row f1 (the update transformer's input rows), produced by the REAL reference's functions here:

    python tests/golden/make_golden_input.py          # needs /root/reference (build container only)

The reference has no function for this step: MVTracker.forward_iteration does it inline
(mvtracker/models/core/mvtracker/mvtracker.py:331-349 embeddings, :374-386 assembly).  This script calls the reference's
own get_3d_embedding / get_3d_sincos_pos_embed_from_grid / get_1d_sincos_pos_embed_from_grid
(mvtracker/models/core/embeddings.py) and joins their outputs with the same torch calls as those lines, on seeded
inputs.  Output: input_a.npz, input_b.npz (inputs + x).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_import  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    ref_import.load()
    from einops import rearrange
    from mvtracker.models.core.embeddings import (get_1d_sincos_pos_embed_from_grid, get_3d_embedding,
                                                  get_3d_sincos_pos_embed_from_grid)
    for name, S, N, C, Lc, seed in (("a", 12, 5, 128, 256, 5), ("b", 5, 8, 128, 256, 6)):
        g = torch.Generator().manual_seed(seed)
        B, E = 1, 64
        D = (E + 1) * 3 + Lc + C + 2                                           # mvtracker.py:130-149
        start = torch.randn(B, 1, N, 3, generator=g) * 1.5
        coords = start + torch.cumsum(0.02 * torch.randn(B, S, N, 3, generator=g), dim=1)
        coords[:, 0] = start[:, 0]
        fcorrs = torch.randn(B, S, N, Lc, generator=g)
        ffeats = torch.randn(B, S, N, C, generator=g)
        track_mask = torch.rand(B, S, N, 1, generator=g) > 0.3
        vis_init = torch.rand(B, S, N, 1, generator=g)
        # mvtracker.py:331-347
        embed_dim = D
        if embed_dim % 6 != 0:
            embed_dim += 6 - (embed_dim % 6)
        pos_embed = get_3d_sincos_pos_embed_from_grid(embed_dim, coords[:, 0:1]).float()[:, 0].permute(0, 2, 1)
        if embed_dim > D:
            pos_embed = pos_embed[:, :D, :]
        pos_embed = rearrange(pos_embed, "b e n -> (b n) e").unsqueeze(1)
        times_ = torch.linspace(0, S - 1, S).reshape(1, S, 1) / S
        embed_dim = D
        if embed_dim % 2 != 0:
            embed_dim += 2 - (embed_dim % 2)
        times_embed = torch.from_numpy(get_1d_sincos_pos_embed_from_grid(embed_dim, times_[0]))[None].repeat(B, 1, 1).float()
        if embed_dim > D:
            times_embed = times_embed[:, :, :D]
        # mvtracker.py:349, 374-386
        track_mask_and_vis = torch.cat([track_mask, vis_init], dim=3).permute(0, 2, 1, 3).reshape(B * N, S, 2)
        fcorrs_ = fcorrs.permute(0, 2, 1, 3).reshape(B * N, S, Lc)
        flows_ = (coords - coords[:, 0:1]).permute(0, 2, 1, 3).reshape(B * N, S, 3)
        flows_ = get_3d_embedding(flows_, E, cat_coords=True)
        ffeats_ = ffeats.permute(0, 2, 1, 3).reshape(B * N, S, C)
        transformer_input = torch.cat([flows_, fcorrs_, ffeats_, track_mask_and_vis], dim=2)
        assert transformer_input.shape[-1] == pos_embed.shape[-1] == D
        x = transformer_input + pos_embed + times_embed
        np.savez_compressed(os.path.join(OUT, f"input_{name}.npz"), coords=coords[0].numpy(), fcorrs=fcorrs[0].numpy(),
                            ffeats=ffeats[0].numpy(), track_mask=track_mask[0, :, :, 0].numpy(), vis=vis_init[0, :, :, 0].numpy(),
                            pos_embed=pos_embed[:, 0].numpy(), time_embed=times_embed[0].numpy(), x=x.numpy(),
                            x_noembed=transformer_input.numpy(), flow_embed_dim=np.int32(E))
        print("input", name, tuple(x.shape))


if __name__ == "__main__":
    main()
