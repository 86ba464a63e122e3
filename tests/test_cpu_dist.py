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


# ---- mode "frames" (dist.FrameShardedSequence): frame ownership, the per-rank cloud cache, window positions, the exchange ----------
class _FakeFrameCloud:
    def __init__(self, level, frames, cam):
        self.level, self.frames, self.cam, self.fvec = level, frames, cam, None


class FrameBackend:
    """Stand-in with the arena / kernel interface FrameShardedSequence uses.  A frame's "cloud" is what was built into its slot;
    the "kernel" writes, for every local frame, a value that depends on that frame's inputs, its camera and its window
    position's targets / coords -- any mix-up of slots, positions or cameras changes it."""

    class CloudArena:
        def __init__(self, V, H, W, C, dtype, device, levels, slots, stride, want_valid=False):
            self.slots, self.levels, self.store, self.moves, self.built = slots, list(levels), {}, 0, []

        def build(self, slot0, fmaps, depths, cam_run, cam_view_stride):
            V, n = fmaps.shape[:2]
            for i in range(n):
                assert 0 <= slot0 + i < self.slots
                cam_i = torch.stack([cam_run[(v * cam_view_stride + i) * 4:(v * cam_view_stride + i) * 4 + 4] for v in range(V)])
                self.store[slot0 + i] = (fmaps[:, i].clone(), depths[:, i].clone(), cam_i.clone())
                self.built.append(slot0 + i)

        def run(self, slot0, n, cam):
            assert slot0 + n <= self.slots
            return [_FakeFrameCloud(l, [self.store[slot0 + i] for i in range(n)], cam) for l in self.levels]

        def move(self, src, dst, n):
            assert src >= dst + n
            for i in range(n):
                self.store[dst + i] = self.store[src + i]
            self.moves += 1

    @staticmethod
    def camera_prep(intrs, extrs):          # [V, n, 3, 3], [V, n, 3, 4] -> [V * n, 4] records (v * n + i)
        V, n = intrs.shape[:2]
        return torch.cat([intrs.reshape(V * n, 9)[:, :1], extrs.reshape(V * n, 12)[:, 9:12]], 1).contiguous()

    @staticmethod
    def plain_peers(t):
        return t

    @staticmethod
    def frame_value(level, fm, dp, cam, targets_f, coords_f, k):
        a = fm[..., ::1 << level, ::1 << level].double().sum() + dp.double().sum() + (cam.double() * torch.arange(1, 5)).sum()
        return (a + targets_f.double().sum(-1, keepdim=True) * (level + 1) + coords_f.double().sum(-1, keepdim=True)).float().expand(-1, k * 4)

    @classmethod
    def knn_corr_shared(cls, levels, targets, coords, frame0, frame_step, out, width):
        for li, (c, _, k) in enumerate(levels):
            for i, (fm, dp, cam_built) in enumerate(c.frames):
                V = fm.shape[0]
                cam_now = torch.stack([c.cam[v * len(c.frames) + i] for v in range(V)])
                assert torch.equal(cam_now, cam_built)               # the window's record of this frame = the one it was built with
                f = frame0 + i * frame_step
                out[f, :, li * k * 4:(li + 1) * k * 4] = cls.frame_value(c.level, fm, dp, cam_now, targets[f], coords[f], k)


def _frames_inputs(T=30, V=2, C=3, H=4, W=8, M=5, seed=3):
    g = torch.Generator().manual_seed(seed)
    fm = torch.randn(V, T, C, H, W, generator=g)
    dp = torch.rand(V, T, H, W, generator=g)
    K = torch.randn(V, T, 3, 3, generator=g)
    E = torch.randn(V, T, 3, 4, generator=g)
    return fm, dp, K, E, M


def _frames_worker(rank, world, port, S, hop, n_win, ret):
    from mvtracker_b200.dist import FrameShardedSequence
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fm, dp, K, E, M = _frames_inputs()
        most = (S + world - 1) // world
        seq = FrameShardedSequence(fm.shape[0], 4, 8, 3, torch.float32, "cpu", S=S, hop=hop, levels=(0, 1), k=2, backend=FrameBackend(),
                                   slots=2 * most)
        assert seq.transport == "allreduce"
        outs = []
        for w in range(n_win):
            win = slice(hop * w, hop * w + S)
            if seq.has_prefetch:
                seq.push()                      # the window handed over by prefetch() during the previous one
            else:
                ts = seq.new_frames()
                assert all(t % world == rank for t in ts)
                seq.push(fm[:, ts] if ts else None, dp[:, ts] if ts else None, K[:, win], E[:, win])
            g = torch.Generator().manual_seed(100 + w)
            targets, coords = torch.randn(S, M, 3, generator=g), torch.randn(S, M, 3, generator=g)
            outs.append(seq.corr(targets, coords).clone().numpy())
            if w % 2 == 0 and w + 1 < n_win:    # every other window is staged ahead (on a CPU device: built at once, adopted by push())
                nxt = slice(hop * (w + 1), hop * (w + 1) + S)
                ts = seq.new_frames()
                seq.prefetch(fm[:, ts] if ts else None, dp[:, ts] if ts else None, K[:, nxt], E[:, nxt])
                with pytest.raises(RuntimeError):
                    seq.prefetch(None, None, K[:, nxt], E[:, nxt])
                outs[-1] = seq.corr(targets, coords).clone().numpy()      # the current window is untouched by the staging
        ret[rank] = (outs, seq.frames_built, sorted(seq.arena.built), seq.arena.moves)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,S,hop,n_win", [(2, 12, 6, 4), (3, 12, 6, 4), (2, 4, 1, 10)])
def test_frame_sharded_sequence_over_gloo(world, S, hop, n_win):
    port = 31500 + os.getpid() % 2000 + 10 * world + S
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_frames_worker, args=(world, port, S, hop, n_win, ret), nprocs=world, join=True)
    fm, dp, K, E, M = _frames_inputs()
    be = FrameBackend
    for w in range(n_win):
        g = torch.Generator().manual_seed(100 + w)
        targets, coords = torch.randn(S, M, 3, generator=g), torch.randn(S, M, 3, generator=g)
        want = torch.empty(S, M, 2 * 2 * 4)
        for f in range(S):
            t = hop * w + f
            cam = be.camera_prep(K[:, t:t + 1], E[:, t:t + 1])
            for l in (0, 1):
                want[f, :, l * 8:(l + 1) * 8] = be.frame_value(l, fm[:, t], dp[:, t], cam, targets[f], coords[f], 2)
        for r in range(world):
            assert np.array_equal(ret[r][0][w], want.numpy()), f"window {w}, rank {r}"
    assert sum(ret[r][1] for r in range(world)) == S + (n_win - 1) * hop        # every frame built once, by its owner
    assert any(ret[r][3] > 0 for r in range(world))                               # the small arenas did wrap around
