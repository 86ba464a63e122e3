"""Drives the REFERENCE's own MVTracker.forward (mvtracker/models/core/mvtracker/mvtracker.py:412-732) on synthetic
sequences, with its stock seam or with mvtracker_b200.dropin.install().

TEST / BASELINE INFRASTRUCTURE ONLY (imported by tests/ and bench.py; never by mvtracker_b200/).  The model code is the
reference's, imported through oracle/ref_import.py (from /root/reference here, from the unmodified copy that
oracle/make_ref.py staged under oracle/_ref on the GPU box).  Weights are random-init (seeded): there is no network for
the checkpoint; what is compared is the SAME model with two implementations of the three seam names.
"""
import time
from types import SimpleNamespace

import numpy as np
import torch

from . import ref_import


def make_sequence(V, T, H, W, N, seed=72, device="cpu", zero_frac=0.05, staggered=True, stride=4):
    """Synthetic multi-view RGB-D sequence of SURVEY.md 8d: ring cameras, analytic depth (ground + dome + moving
    spheres) with a fraction of invalid (0) pixels, uniform random uint8-valued colours; N queries (t, x, y, z) sampled
    from the valid points of their start frame's fused cloud (demo.py:781-806); staggered: half of them start at frames
    spread over [0, T/2) so that later windows introduce queries (mvtracker.py:606-645)."""
    from mvtracker_b200 import synth
    g = torch.Generator().manual_seed(seed)
    K, E = synth.ring_cameras(V, H, W)
    depth = synth.raycast_depth(K, E, T, H, W, 1)                       # [V,T,H,W] at full resolution
    if zero_frac > 0:
        depth = torch.where(torch.rand(depth.shape, generator=g) < zero_frac, torch.zeros_like(depth), depth)
    rgbs = torch.randint(0, 256, (1, V, T, 3, H, W), generator=g).float()
    intrs = K[None, :, None].expand(1, V, T, 3, 3).contiguous()
    extrs = E[None, :, None].expand(1, V, T, 3, 4).contiguous()
    # queries: unproject random valid pixels of a random view at the query's start frame (fp64 on the host)
    ts = torch.zeros(N, dtype=torch.long)
    if staggered:
        ts[N // 2:] = torch.randint(0, max(T // 2, 1), (N - N // 2,), generator=g)
    q = torch.zeros(N, 4)
    K64, E64 = K.double(), E.double()
    for i in range(N):
        while True:
            v = int(torch.randint(0, V, (1,), generator=g))
            y = int(torch.randint(0, H, (1,), generator=g))
            x = int(torch.randint(0, W, (1,), generator=g))
            d = float(depth[v, ts[i], y, x])
            if d > 0:
                break
        cam = torch.linalg.inv(K64[v]) @ torch.tensor([x, y, 1.0], dtype=torch.float64) * d
        R, t = E64[v, :, :3], E64[v, :, 3]
        world = R.T @ (cam - t)
        q[i, 0] = float(ts[i])
        q[i, 1:] = world.float()
    dev = torch.device(device)
    return SimpleNamespace(rgbs=rgbs.to(dev), depths=depth[None, :, :, None].contiguous().to(dev), intrs=intrs.to(dev),
                           extrs=extrs.to(dev), query_points=q[None].to(dev), V=V, T=T, H=H, W=W, N=N)


def build_model(seed=72, device="cpu", lively=0.0, **kw):
    """The reference's MVTracker with seeded random-init weights (demo.py:572-573 seeds 72).
    lively > 0: the update head (EfficientUpdateFormer.flow_head, initialised with std 1e-3 so that an untrained model
    leaves every track where it started, cotracker2/blocks.py:443-451) is re-drawn with Xavier weights and its last layer
    scaled by `lively`: the random model then MOVES tracks and features by centimetres per iteration, so that what the
    seam returns feeds back into where it is asked next -- a parity test of the stock init alone would hardly depend on
    the correlation features."""
    ref = ref_import.load()
    torch.manual_seed(seed)
    np.random.seed(seed)
    cfg = dict(hidden_size=256, use_flash_attention=False)      # configs/model/mvtracker.yaml:13 (flash = SDPA kernels)
    cfg.update(kw)
    model = ref.MVTracker(**cfg)
    if lively > 0:
        g = torch.Generator().manual_seed(seed + 1)
        lin = [m for m in model.updateformer.flow_head if isinstance(m, torch.nn.Linear)]
        for m in lin:
            bound = (6.0 / (m.in_features + m.out_features)) ** 0.5
            with torch.no_grad():
                m.weight.copy_((torch.rand(m.weight.shape, generator=g) * 2 - 1) * bound)
        with torch.no_grad():
            lin[-1].weight.mul_(lively)
    return model.to(device).eval()


def knn_exact(k, xyz_ref, xyz_query):
    """_knn_torch (mvtracker.py:75-79) with cdist's exact-difference mode instead of the |x|^2+|y|^2-2xy expansion
    (the expansion errs by up to 8.7e-3 and flips ~1 % of neighbour sets 10 units from the origin).  Used as
    the adjudicator arm: the reference model, the reference cloud build and correlation, exact distances."""
    d = torch.cdist(xyz_query, xyz_ref, p=2, compute_mode="donot_use_mm_for_euclid_dist")
    return torch.topk(d, k, dim=-1, largest=False, sorted=True)


def forward(model, seq, iters=4, seam="stock", autocast=None):
    """seam: 'stock' (reference as is), 'exact' (reference with knn_exact), 'dropin' (mvtracker_b200.dropin.install()).
    -> (results dict, seconds, seam call counts)"""
    ref = ref_import.load()
    undo = None
    if seam == "dropin":
        from mvtracker_b200 import dropin
        undo = dropin.install(ref.mvt)
    elif seam == "exact":
        saved = ref.mvt.knn
        ref.mvt.knn = knn_exact

        def undo():
            ref.mvt.knn = saved
    elif seam != "stock":
        raise ValueError(seam)
    try:
        dev = seq.rgbs.device
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        with torch.no_grad():
            if autocast is not None:
                with torch.autocast(device_type=dev.type, dtype=autocast):
                    out = model(rgbs=seq.rgbs, depths=seq.depths, query_points=seq.query_points, intrs=seq.intrs,
                                extrs=seq.extrs, iters=iters)
            else:
                out = model(rgbs=seq.rgbs, depths=seq.depths, query_points=seq.query_points, intrs=seq.intrs,
                            extrs=seq.extrs, iters=iters)
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
    finally:
        if undo is not None:
            undo()
    return out, dt


def compare(a, b):
    """Differences of two forward() results.  feat_init[:, :, n] is the feature row of the cloud point nearest to query n
    (k = 1, mvtracker.py:627-645): where two kNN implementations pick different points for a (near-)tie the whole track
    starts from another feature vector, so visibilities are compared on the tracks whose rows agree (`vis`), and over all
    tracks (`vis_all`)."""
    same = (a["feat_init"].float() == b["feat_init"].float()).all(-1).all(1)[0]           # [N]
    dv = (a["vis_e"].float() - b["vis_e"].float()).abs()[0]                                # [T, N]
    dt = (a["traj_e"].float() - b["traj_e"].float()).abs()[0].amax(-1)                     # [T, N]
    return {"traj": float(dt[:, same].max()) if same.any() else 0.0, "traj_all": float(dt.max()),
            "vis": float(dv[:, same].max()) if same.any() else 0.0, "vis_all": float(dv.max()),
            "feat_rows_equal": float(same.float().mean()), "scale": float(a["traj_e"].float().abs().max())}
