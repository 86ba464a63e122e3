"""ctypes/numpy front-end of the CPU ORACLE (oracle/mvt_oracle.c).

TEST INFRASTRUCTURE ONLY.  May be imported by tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference leg -- never by mvtracker_b200/.

Each function restates one reference call (paths under /root/reference):
  unproject    -> mvtracker/models/core/model_utils.py:420-482  init_pointcloud_from_rgbd
  knn          -> mvtracker/models/core/mvtracker/mvtracker.py:75-79 (canonical exact-fp32 arithmetic,
                  (d2, index) order; see the header of mvt_oracle.c)
  corr_sample  -> mvtracker/models/core/mvtracker/mvtracker.py:800-846
Arrays are numpy; B (the reference's leading batch axis, asserted == 1 at
mvtracker.py:275) is squeezed: fmaps [V,S,C,H,W], depths [V,S,H,W].
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libmvt_oracle.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "mvt_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        for name in ("orc_unproject", "orc_knn", "orc_corr_from_idx", "orc_corr_sample"):
            getattr(_lib, name).restype = C.c_int
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def num_threads():
    return int(lib().orc_num_threads())


def set_num_threads(n):
    lib().orc_set_num_threads(C.c_int(int(n)))


def unproject(fmaps, depths, intrs, extrs, stride=4, level=0, depth_interp_mode="nearest", return_validity_mask=False):
    fmaps, depths, intrs, extrs = _f32(fmaps), _f32(depths), _f32(intrs), _f32(extrs)
    V, S, Cc, H, W = fmaps.shape
    assert depths.shape == (V, S, H, W) and intrs.shape == (V, S, 3, 3) and extrs.shape == (V, S, 3, 4)
    Hl, Wl = H >> level, W >> level
    P = V * Hl * Wl
    xyz = np.empty((S, P, 3), np.float32)
    fvec = np.empty((S, P, Cc), np.float32)
    valid = np.empty((S, P), np.uint8)
    rc = lib().orc_unproject(_p(fmaps), _p(depths), _p(intrs), _p(extrs), V, S, Cc, H, W, C.c_float(stride),
                             int(level), int(depth_interp_mode == "avg"), _p(xyz), _p(fvec), _p(valid))
    if rc:
        raise RuntimeError(f"orc_unproject failed rc={rc}")
    if return_validity_mask:
        return xyz, fvec, valid.astype(bool)
    return xyz, fvec


def knn(k, xyz_ref, xyz_query, mask=None):
    """-> (dists [B,M,k] f32 ascending, idx [B,M,k] int32), canonical order."""
    ref, qry = _f32(xyz_ref), _f32(xyz_query)
    B, N, _ = ref.shape
    M = qry.shape[1]
    assert qry.shape == (B, M, 3)
    mk = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
    idx = np.empty((B, M, k), np.int32)
    dist = np.empty((B, M, k), np.float32)
    rc = lib().orc_knn(int(k), _p(ref), _p(mk), _p(qry), B, N, M, _p(idx), _p(dist))
    if rc:
        raise RuntimeError(f"orc_knn failed rc={rc}")
    return dist, idx


def corr_from_idx(xyz, fvec, targets, coords, idx, groups=1, add_offset=True, add_xyz=False):
    xyz, fvec, targets, coords = _f32(xyz), _f32(fvec), _f32(targets), _f32(coords)
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    B, N, Cc = fvec.shape
    M, k = idx.shape[1], idx.shape[2]
    D = groups + 3 * bool(add_offset) + 3 * bool(add_xyz)
    out = np.empty((B, M, k, D), np.float32)
    rc = lib().orc_corr_from_idx(_p(xyz), _p(fvec), _p(targets), _p(coords), _p(idx), B, N, M, Cc, k, int(groups),
                                 int(bool(add_offset)), int(bool(add_xyz)), _p(out))
    if rc:
        raise RuntimeError(f"orc_corr_from_idx failed rc={rc}")
    return out


def corr_sample(k, xyz, fvec, targets, coords, groups=1, add_offset=True, add_xyz=False, mask=None,
                return_idx=False):
    """One PointcloudCorrBlock.corr_sample: kNN -> gather -> grouped dot / sqrt(C/G) -> offsets."""
    dist, idx = knn(k, xyz, coords, mask)
    out = corr_from_idx(xyz, fvec, targets, coords, idx, groups, add_offset, add_xyz)
    return (out, idx, dist) if return_idx else out


def dist2_f64(ref, q):
    ref = _f32(ref)
    q = _f32(q)
    out = np.empty((ref.shape[0],), np.float64)
    lib().orc_dist2_f64(_p(ref), _p(q), ref.shape[0], _p(out))
    return out


def transformer_input(coords, fcorrs, ffeats, track_mask, vis, pos_embed=None, time_embed=None, flow_embed_dim=64):
    """numpy restatement of the assembly of the update transformer's input:
    mvtracker/models/core/mvtracker/mvtracker.py:374-386 with get_3d_embedding (mvtracker/models/core/embeddings.py:134-161).
    coords [S,N,3], fcorrs [S,N,Lc], ffeats [S,N,C], track_mask / vis [S,N], pos_embed [N,D], time_embed [S,D]
    -> x [N,S,D] fp32, D = 3E+3 + Lc + C + 2.  All arithmetic in fp32, in the reference's order."""
    f32 = np.float32
    coords = np.asarray(coords, f32)
    S, N, _ = coords.shape
    E = int(flow_embed_dim)
    flows = (coords - coords[0:1]).transpose(1, 0, 2)                                  # [N,S,3]  (:379)
    div = np.arange(0, E, 2, dtype=f32) * f32(1000.0 / E)                              # embeddings.py:141-143
    pe = np.zeros((N, S, 3, E), f32)
    arg = flows[..., None] * div[None, None, None, :]                                  # fp32 products
    pe[..., 0::2] = np.sin(arg.astype(np.float64)).astype(f32)                         # correctly rounded sin / cos of the fp32 argument
    pe[..., 1::2] = np.cos(arg.astype(np.float64)).astype(f32)
    emb = np.concatenate([pe.reshape(N, S, 3 * E), flows], axis=2)                     # cat_coords=True: pe first (:159-160)
    tm = np.asarray(track_mask).reshape(S, N).astype(f32).T[..., None]
    vs = np.asarray(vis, f32).reshape(S, N).T[..., None]
    x = np.concatenate([emb, np.asarray(fcorrs, f32).transpose(1, 0, 2), np.asarray(ffeats).astype(f32).transpose(1, 0, 2), tm, vs],
                       axis=2)                                                         # (:384)
    if pos_embed is not None:
        x = x + np.asarray(pos_embed, f32).reshape(N, 1, -1)                           # (:386) x = (input + pos) + time
    if time_embed is not None:
        x = x + np.asarray(time_embed, f32).reshape(1, S, -1)
    return x.astype(f32)


def corr_backward(fvec, targets, idx, grad_out, groups=1):
    """numpy restatement of what autograd computes through corr_sample's gather + einsum / sqrt(C/G)
    (mvtracker/models/core/mvtracker/mvtracker.py:827-835) given the neighbour lists.
    fvec [B,N,C], targets [B,M,C], idx [B,M,k] (-1 = none), grad_out [B,M,k,D] -> (grad_targets [B,M,C], grad_fvec [B,N,C]),
    accumulated in float64 and rounded once (the CUDA reductions add in an unspecified order)."""
    fvec, targets, g = np.asarray(fvec, np.float64), np.asarray(targets, np.float64), np.asarray(grad_out, np.float64)
    B, N, Cc = fvec.shape
    M, k = idx.shape[1], idx.shape[2]
    G = int(groups)
    Cg = Cc // G
    gs = g[..., :G] / np.float64(np.sqrt(np.float32(Cg)))                     # [B,M,k,G]
    gs = np.repeat(gs, Cg, axis=-1)                                           # per channel [B,M,k,C]
    gt = np.zeros((B, M, Cc))
    gf = np.zeros((B, N, Cc))
    for b in range(B):
        for j in range(k):
            ij = idx[b, :, j]
            ok = ij >= 0
            gt[b, ok] += gs[b, ok, j] * fvec[b, ij[ok]]
            np.add.at(gf[b], ij[ok], gs[b, ok, j] * targets[b, ok])
    return gt.astype(np.float32), gf.astype(np.float32)


def unproject_backward(grad_fvec, V, H, W, level):
    """Feature branch of init_pointcloud_from_rgbd backwards (mvtracker/models/core/model_utils.py:439-441 chained
    avg_pool2d(2, 2), :475-476 permute + reshape): grad_fvec [S, V*Hl*Wl, C] -> grad_fmaps [V,S,C,H,W]."""
    g = np.asarray(grad_fvec, np.float32)
    S, P, Cc = g.shape
    Hl, Wl = H >> level, W >> level
    assert P == V * Hl * Wl
    gl = g.reshape(S, V, Hl, Wl, Cc).transpose(1, 0, 4, 2, 3) * np.float32(1.0 / (1 << (2 * level)))     # [V,S,C,Hl,Wl]
    out = np.zeros((V, S, Cc, H, W), np.float32)
    f = 1 << level
    out[:, :, :, :Hl * f, :Wl * f] = np.repeat(np.repeat(gl, f, axis=3), f, axis=4)
    return out


def lift_pixels(depth, K, E, pixels_xy):
    """numpy restatement of the support-grid unprojection: bilinear_sample2d of the depth
    (mvtracker/models/core/model_utils.py:81-165) and camera = Kinv [x,y,1] z, world = Einv [camera,1]
    (mvtracker/models/evaluation_predictor_3dpt.py:103-116).  depth [H,W], K [3,3], E [3,4], pixels_xy [n,2] -> [n,3]."""
    f32 = np.float32
    d = np.asarray(depth, f32)
    H, W = d.shape
    x, y = np.asarray(pixels_xy, f32)[:, 0], np.asarray(pixels_xy, f32)[:, 1]
    x0, y0 = np.floor(x), np.floor(y)
    x1, y1 = x0 + 1, y0 + 1
    xi0, xi1 = np.clip(x0.astype(np.int64), 0, W - 1), np.clip(x1.astype(np.int64), 0, W - 1)
    yi0, yi1 = np.clip(y0.astype(np.int64), 0, H - 1), np.clip(y1.astype(np.int64), 0, H - 1)
    z = ((x1 - x) * (y1 - y)) * d[yi0, xi0] + ((x - x0) * (y1 - y)) * d[yi0, xi1] + ((x1 - x) * (y - y0)) * d[yi1, xi0] \
        + ((x - x0) * (y - y0)) * d[yi1, xi1]
    Kinv = np.linalg.inv(np.asarray(K, np.float64))
    E4 = np.eye(4)
    E4[:3] = np.asarray(E, np.float64)
    Einv = np.linalg.inv(E4)
    cam = (Kinv @ np.stack([x, y, np.ones_like(x)]).astype(np.float64)) * z.astype(np.float64)
    world = Einv @ np.concatenate([cam, np.ones((1, cam.shape[1]))])
    return (world[:3] / world[3:]).T.astype(f32)
