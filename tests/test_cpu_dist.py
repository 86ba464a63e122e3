"""Host-side logic of the multi-GPU path (mvtracker_b200/dist.py) on CPU: world_size-2/3 gloo process groups drive the
real exchange code with a numpy stand-in for the three CUDA kernels (same layouts and semantics as include/mvt.h)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mvtracker_b200.dist import ShardedWindow, track_shard, view_range  # noqa: E402

EMPTY = 0x7F8000007FFFFFFF


class FakeCloud:
    def __init__(self, xyz, fvec, H, W):
        self.xyz4, self.fvec, self.H, self.W = xyz, fvec, H, W     # xyz4 [S, Nloc, 4] (w unused)


class NumpyBackend:
    """CPU stand-in with the C ABI's layouts: packed keys (bits(d2) << 32 | global index), -1 for missing indices."""
    EMPTY_KEY = EMPTY

    @staticmethod
    def camera_prep(intrs, extrs):
        return None

    @staticmethod
    def full_cloud(level, xyz4, cam, V, H, W, stride):
        return FakeCloud(xyz4, None, H, W)

    @staticmethod
    def _d2(p, q):
        d = p.astype(np.float32) - q.astype(np.float32)
        return ((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]).astype(np.float32)

    def knn_keys(self, clouds, offs, coords, k, n_shards, out, caps=None, budget=0, exact_out=None, skip=None):
        # budget: "gave up" is emulated for every third query, which then returns every other point of the exact list
        # (k real points = a valid bound) and must be redone in the second phase
        S, M, _ = coords.shape
        Ms = track_shard(n_shards, M)
        if skip is None:
            out.fill_(EMPTY)
        o = out.numpy()
        sk = None if skip is None else skip.numpy().reshape(n_shards, -1, S, Ms)
        ex = None if exact_out is None else exact_out.numpy().reshape(n_shards, -1, S, Ms)
        q = coords.numpy()
        for l, (c, off) in enumerate(zip(clouds, offs)):
            pts = c.xyz4.numpy()[..., :3]
            for s in range(S):
                d2 = self._d2(pts[s][None, :, :], q[s][:, None, :])                     # [M, Nloc]
                keys = (d2.view(np.uint32).astype(np.int64) << 32) | (np.arange(pts.shape[1], dtype=np.int64) + off)[None]
                keys.sort(axis=1)
                kk = min(k, keys.shape[1])
                for m in range(M):
                    if sk is not None and sk[m // Ms, l, s, m % Ms]:
                        continue
                    o[m // Ms, l, s, m % Ms, :] = EMPTY
                    row = keys[m, :kk]
                    gave_up = budget and (m + s + l) % 3 == 0
                    if ex is not None:
                        ex[m // Ms, l, s, m % Ms] = 0 if gave_up else 1
                    if gave_up:
                        row = keys[m, 0:2 * kk:2]
                    if caps is not None:
                        cap = int(caps.view(n_shards, -1, S, Ms)[m // Ms, l, s, m % Ms])
                        row = row[(row >> 32) <= (cap >> 32)]
                    o[m // Ms, l, s, m % Ms, :len(row)] = row
        return out

    def topk_merge(self, keys, out=None, keys_out=None):
        a = keys.numpy()                               # [n_lists, Q, k]
        n, Q, k = a.shape
        merged = np.sort(a.transpose(1, 0, 2).reshape(Q, n * k), axis=1)[:, :k]
        if keys_out is not None:
            keys_out.copy_(torch.from_numpy(merged.copy()))
            return keys_out
        idx = (merged & 0xFFFFFFFF).astype(np.int64)
        idx[merged == EMPTY] = -1
        out.copy_(torch.from_numpy(idx.astype(np.int32)))
        return out

    def corr_from_idx(self, owned, idx, targets, coords, k, out):
        S, M, C = targets.shape
        n_shards, L = idx.shape[0], idx.shape[1]
        Ms = idx.shape[3]
        o = out.numpy()
        o[...] = 0
        ii, tg, q = idx.numpy(), targets.numpy(), coords.numpy()
        for l, (xyz4, fvec, lo, hi) in enumerate(owned):
            P, F = xyz4.numpy(), fvec.numpy()
            for s in range(S):
                for m in range(M):
                    g = ii[m // Ms, l, s, m % Ms]
                    for j in range(k):
                        if lo <= g[j] < hi:
                            r = g[j] - lo
                            rec = o[s, m, l * k * 4 + j * 4: l * k * 4 + j * 4 + 4]
                            rec[0] = np.float32(np.dot(tg[s, m], F[s, r]) / np.sqrt(np.float32(C)))
                            rec[1:] = P[s, r, :3] - q[s, m]
        return out


def _scene(seed=5, V=5, S=2, Hs=(6, 3), Ws=(8, 4), C=8, M=11):
    g = np.random.default_rng(seed)
    levels = []
    for H, W in zip(Hs, Ws):
        xyz = g.normal(size=(S, V * H * W, 4)).astype(np.float32)
        xyz[:, ::7] = xyz[:, 1:2]                       # exact duplicates: ties must break by global index
        levels.append((xyz, g.normal(size=(S, V * H * W, C)).astype(np.float32), H, W))
    coords = g.normal(size=(S, M, 3)).astype(np.float32)
    targets = g.normal(size=(S, M, C)).astype(np.float32)
    return levels, coords, targets, V


def _worker(rank, world, port, k, V, mode, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        levels, coords, targets, V = _scene(V=V)
        v0, v1 = view_range(rank, world, V)
        clouds = [FakeCloud(torch.from_numpy(x[:, v0 * H * W:v1 * H * W].copy()), torch.from_numpy(f[:, v0 * H * W:v1 * H * W].copy()), H, W)
                  for x, f, H, W in levels] if v1 > v0 else []
        win = ShardedWindow(clouds, v0, v1, V, k=k, backend=NumpyBackend(), n_levels=len(levels))
        if mode == "tracks":
            win.mode = "tracks"
            win._replicate_records(torch.empty(v1 - v0, coords.shape[0], 1, levels[0][2], levels[0][3]), None, None, 4, [0, 1])
        out = win.corr(torch.from_numpy(targets), torch.from_numpy(coords))
        ret[rank] = out.numpy().copy()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["tracks", "views"])
@pytest.mark.parametrize("world,V", [(2, 4), (3, 5), (3, 2)])      # (3, 5): uneven shards; (3, 2): a rank without views
def test_sharded_window_over_gloo(world, V, mode):
    k = 4
    port = 29500 + os.getpid() % 2000 + 10 * world + V + (100 if mode == "tracks" else 0)
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, k, V, mode, ret), nprocs=world, join=True)
    # single-process answer with the same stand-in arithmetic
    levels, coords, targets, V = _scene(V=V)
    be = NumpyBackend()
    clouds = [FakeCloud(torch.from_numpy(x), torch.from_numpy(f), H, W) for x, f, H, W in levels]
    S, M, _ = targets.shape
    keys = torch.empty(1, len(levels), S, M, k, dtype=torch.int64)
    be.knn_keys(clouds, [0] * len(levels), torch.from_numpy(coords), k, 1, keys)
    idx = torch.empty(len(levels) * S * M, k, dtype=torch.int32)
    be.topk_merge(keys.view(1, -1, k), out=idx)
    want = torch.empty(S, M, len(levels) * k * 4)
    be.corr_from_idx([(c.xyz4, c.fvec, 0, c.xyz4.shape[1]) for c in clouds], idx.view(1, len(levels), S, M, k),
                     torch.from_numpy(targets), torch.from_numpy(coords), k, want)
    for r in range(world):
        assert np.array_equal(ret[r], want.numpy()), f"rank {r} differs from the single-process result"
    assert np.abs(want.numpy()).sum() > 0


def test_view_ranges_partition_the_views():
    for V in (1, 3, 8, 16):
        for world in (1, 2, 3, 4, 8):
            spans = [view_range(r, world, V) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == V
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(b - a for a, b in spans) - min(b - a for a, b in spans) <= 1
    assert track_shard(8, 4096) == 512 and track_shard(3, 10) == 4
