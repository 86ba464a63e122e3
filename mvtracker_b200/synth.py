"""Synthetic multi-view RGB-D scenes of the shapes BASELINE.json names.

No network / datasets: cameras sit on a ring (radius 4 m, height 1 m) looking at the
origin, depth is an analytic ray-cast of a ground plane, a bounding dome and a few
moving spheres so the per-view clouds overlap, with an optional fraction of pixels set
to 0 ("DROID-like" invalid depth, conversions/droid/debug/generate_npz_from_external_cams.py:231).
Everything is seeded; tensors are produced on the requested device.
"""
import math
from dataclasses import dataclass

import torch

# name -> (V, H, W, N tracks, T frames)   (H,W are input resolution, stride 4)
WORKLOADS = {
    "kubric": dict(V=4, H=256, W=256, N=512, T=24),          # config 1
    "droid": dict(V=3, H=720, W=1280, N=2048, T=64),         # config 2
    "panoptic": dict(V=8, H=1280, W=1600, N=4096, T=150),    # config 3 (~1M fused points / frame)
    "dexycb": dict(V=8, H=480, W=640, N=16384, T=100),       # config 4
    "tiny": dict(V=2, H=64, W=96, N=64, T=12),               # smoke / unit tests
}


@dataclass
class Scene:
    fmaps: torch.Tensor     # [1,V,S,C,H',W']
    depths: torch.Tensor    # [1,V,S,1,H',W']
    intrs: torch.Tensor     # [1,V,S,3,3]
    extrs: torch.Tensor     # [1,V,S,3,4]  world -> camera
    stride: int


def ring_cameras(V, H, W, device="cpu", radius=4.0, height=1.0):
    """-> K [V,3,3], E [V,3,4] (world->cam), f = 0.86*W, principal point at the centre."""
    K = torch.zeros(V, 3, 3, dtype=torch.float64)
    E = torch.zeros(V, 3, 4, dtype=torch.float64)
    for v in range(V):
        th = 2.0 * math.pi * v / V + 0.1
        pos = torch.tensor([radius * math.cos(th), radius * math.sin(th), height], dtype=torch.float64)
        zc = -pos / pos.norm()
        up = torch.tensor([0.0, 0.0, 1.0], dtype=torch.float64)
        xc = torch.linalg.cross(zc, up)
        xc = xc / xc.norm()
        yc = torch.linalg.cross(zc, xc)
        R = torch.stack([xc, yc, zc])
        E[v, :, :3] = R
        E[v, :, 3] = -R @ pos
        f = 0.86 * W
        K[v] = torch.tensor([[f, 0, W / 2.0], [0, f, H / 2.0], [0, 0, 1.0]], dtype=torch.float64)
    return K.float().to(device), E.float().to(device)


def raycast_depth(K, E, S, Hs, Ws, stride, t0=0, device="cpu"):
    """z-depth at the nearest-strided pixel centres (stride*i, stride*j): [V,S,Hs,Ws] fp32."""
    V = K.shape[0]
    K64, E64 = K.double().cpu(), E.double().cpu()
    ys = (torch.arange(Hs, dtype=torch.float64) * stride)
    xs = (torch.arange(Ws, dtype=torch.float64) * stride)
    out = torch.empty(V, S, Hs, Ws, dtype=torch.float32, device=device)
    for v in range(V):
        Kinv = torch.linalg.inv(K64[v])
        R, t = E64[v, :, :3], E64[v, :, 3]
        o = (-R.T @ t).to(device)                                   # camera centre in world
        pix = torch.stack([xs[None, :].expand(Hs, Ws), ys[:, None].expand(Hs, Ws), torch.ones(Hs, Ws, dtype=torch.float64)], -1)
        d = ((pix @ Kinv.T) @ R).to(device)                          # world dir per unit camera z  (R^T r)
        for s in range(S):
            tt = (t0 + s) * 0.05
            best = torch.full((Hs, Ws), float("inf"), dtype=torch.float64, device=device)
            # ground plane z = 0
            tz = -o[2] / d[..., 2]
            best = torch.where((tz > 0.05) & (tz < best), tz, best)
            # bounding dome (inside hit of a sphere radius 6 at the origin) and moving spheres
            spheres = [((0.0, 0.0, 0.0), 6.0, True)]
            for j in range(4):
                c = (1.2 * math.cos(1.7 * j + tt), 1.2 * math.sin(1.7 * j + 0.7 * tt), 0.45 + 0.1 * j)
                spheres.append((c, 0.35 + 0.05 * j, False))
            for c, r, inside in spheres:
                c_t = torch.tensor(c, dtype=torch.float64, device=device)
                oc = o - c_t
                a = (d * d).sum(-1)
                b = 2.0 * (d * oc).sum(-1)
                cc = (oc * oc).sum() - r * r
                disc = b * b - 4 * a * cc
                sq = torch.sqrt(disc.clamp_min(0))
                th = ((-b + sq) if inside else (-b - sq)) / (2 * a)
                ok = (disc > 0) & (th > 0.05) & (th < best)
                best = torch.where(ok, th, best)
            best = torch.where(torch.isfinite(best), best, torch.zeros_like(best))
            out[v, s] = best.float()
    return out


def make_scene(V, H, W, S=12, C=128, stride=4, zero_frac=0.05, seed=72, device="cpu", dtype=torch.float32, t0=0):
    """Hot-path inputs of one sliding window: fmaps ~ N(0,1) (the encoder is out of scope), analytic depth."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    Hs, Ws = H // stride, W // stride
    K, E = ring_cameras(V, H, W, device)
    depth = raycast_depth(K, E, S, Hs, Ws, stride, t0, device)
    if zero_frac > 0:
        drop = torch.rand(depth.shape, generator=g, device=device) < zero_frac
        depth = torch.where(drop, torch.zeros_like(depth), depth)
    fmaps = torch.randn(1, V, S, C, Hs, Ws, generator=g, device=device, dtype=torch.float32).to(dtype)
    intrs = K[None, :, None].expand(1, V, S, 3, 3).contiguous()
    extrs = E[None, :, None].expand(1, V, S, 3, 4).contiguous()
    return Scene(fmaps=fmaps, depths=depth[None, :, :, None].contiguous(), intrs=intrs, extrs=extrs, stride=stride)


def make_tracks(xyz0, N, S, C=128, jitter=0.02, seed=73, valid=None, dtype=torch.float32):
    """Tracks as in demo.py:781-806: N points of the fused level-0 cloud, per-frame coords = point + N(0, jitter).
    xyz0 [P,3]: frame 0's cloud, the same point in every frame (static tracks);
    xyz0 [S,P,3]: the track follows the surface seen at its pixel in every frame (stays near the cloud even when the
    scene moves, as converged tracks do).  valid [P] / [S,P] optional.  -> coords [S,N,3] fp32, targets [S,N,C]."""
    device = xyz0.device
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    per_frame = xyz0.dim() == 3
    P = xyz0.shape[-2]
    pool = torch.arange(P, device=device)
    if valid is not None:
        v = valid.bool()
        pool = pool[v.all(0) if v.dim() == 2 else v]
    sel = pool[torch.randint(0, pool.numel(), (N,), generator=g, device=device)]
    base = xyz0[:, sel].float() if per_frame else xyz0[sel].float()[None]
    coords = base + jitter * torch.randn(S, N, 3, generator=g, device=device)
    targets = torch.randn(S, N, C, generator=g, device=device, dtype=torch.float32).to(dtype)
    return coords.contiguous(), targets.contiguous()
