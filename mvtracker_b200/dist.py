"""Multi-GPU hot path: views are sharded for the cloud build, candidates are merged per track shard.

One process per GPU (torch.distributed, NCCL over NVLink).  Rank r unprojects a contiguous range of views and KEEPS
those clouds -- feature rows (512 B per point) never cross NVLink.  Per refinement iteration

    keys   = knn_keys(local clouds, budget)                  K0   bounded-effort local search: exact where it converges
    caps   = all_gather(kth(topk_merge(all_to_all(keys))))   C0   an upper bound of the GLOBAL k-th distance, on every rank
    keys   = knn_keys(local clouds, caps, skip=converged)    K1   [N, L, S, Ms, k] packed (d2, global index), exact within caps
    recv   = all_to_all(keys)                                C1   rank j gets every rank's lists for track shard j
    idx_j  = topk_merge(recv)                                K2   exact global top-k for shard j
    idx    = all_gather(idx_j)                               C2
    out    = corr_from_idx(local clouds, idx)                K3   records of neighbours in local views, zeros elsewhere
    out    = all_reduce(out, SUM)                            C3   exactly one non-zero contributor per record

(without K0/C0 a rank whose views do not see a query's surface searches out to its own k-th distance, metres away:
measured 11.8 ms instead of 0.4 ms per iteration at 2 GPUs) and every rank ends with the full [S, M, L*k*4] block (the transformer that consumes it is replicated and needs all
tracks: reference mvtracker/models/core/cotracker2/blocks.py:473-489).  Point indices are the reference's global ones
(v*H*W + y*W + x, model_utils.py:475), so results equal the single-GPU kernel's bit for bit in the neighbour sets and
offsets; correlations are the same fp32 sums.

north_star's literal plan (all-gather the clouds, then shard tracks) moves S*P*(C*e+16) bytes per window -- 8.6 GB
for the 8-view / 1M-point window -- which at NVLink speed costs as much as the whole single-GPU step; this design
moves ~1000x less ("Multi-GPU").  `backend` is the module providing the three kernels (mvtracker_b200.ops);
the CPU tests inject a numpy stand-in to exercise the exchanges over gloo.
"""
from typing import List, Optional, Sequence, Tuple

import os

import torch
import torch.distributed as dist


class SharedBuffer:
    """A tensor in symmetric memory (same size on every rank of the node) + the addresses the kernels store through:
    the NVSwitch multicast address when the allocation has one, and every rank's peer mapping."""

    def __init__(self, shape, dtype, device, group=None):
        import torch.distributed._symmetric_memory as symm
        from ._lib import Peers
        self.t = symm.empty(*shape, dtype=dtype, device=device)
        self.h = symm.rendezvous(self.t, dist.group.WORLD if group is None else group)
        world, rank = self.h.world_size, self.h.rank
        ptrs = list(self.h.buffer_ptrs)
        off = self.t.data_ptr() - ptrs[rank]
        self.peers = Peers()
        self.peers.n = world
        for r in range(world):
            self.peers.ptr[r] = ptrs[r] + off
        mc = int(getattr(self.h, "multicast_ptr", 0) or 0)
        self.multicast = bool(mc) and not os.environ.get("MVT_NO_MULTICAST")
        self.peers.mc = (mc + off) if self.multicast else None

    def barrier(self):
        """Cross-rank barrier on the current stream (signal pads): everything stored before it, by any rank, is visible after."""
        self.h.barrier(channel=0)


def symmetric_memory_available() -> bool:
    try:
        import torch.distributed._symmetric_memory as symm  # noqa: F401
        return dist.is_initialized() and dist.get_backend() == "nccl"
    except Exception:
        return False


def _all_to_all(recv: torch.Tensor, send: torch.Tensor, group=None):
    """Equal-split all-to-all along dim 0.  NCCL: all_to_all_single.  gloo has no all-to-all (CPU tests only):
    all-gather and slice -- same result, plumbing only."""
    if dist.get_backend(group) != "gloo":
        dist.all_to_all_single(recv, send, group=group)
        return
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    everything = [torch.empty_like(send) for _ in range(world)]
    dist.all_gather(everything, send.contiguous(), group=group)
    for src in range(world):
        recv[src].copy_(everything[src][rank])


def view_range(rank: int, world: int, V: int) -> Tuple[int, int]:
    """Contiguous views [v0, v1) of `rank`; the first V % world ranks get one more.  Ranks beyond V get none."""
    base, rem = divmod(V, world)
    v0 = rank * base + min(rank, rem)
    return v0, v0 + base + (1 if rank < rem else 0)


def track_shard(world: int, M: int) -> int:
    """Tracks per shard (the last shard may be padded)."""
    return (M + world - 1) // world


class ShardedWindow:
    """One sliding window's clouds of THIS rank's views + the exchanges that make corr() global."""

    def __init__(self, clouds, v0: int, v1: int, V_total: int, k: int = 16, group=None, backend=None, n_levels=None):
        if backend is None:
            from . import ops as backend
        self.be, self.group = backend, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.clouds, self.v0, self.v1, self.V_total, self.k = clouds, v0, v1, V_total, k
        self.levels = len(clouds) if n_levels is None else n_levels     # a rank without views still has to know
        self.width = self.levels * k * 4
        self._bufs = {}
        self.split_records = False   # "p2p": only correlation values cross NVLink, offsets are filled in locally
        # phase-A effort per query: the fixed windows (30 pixels per view) plus about one growth step
        self.budget = 4 * 30 * max(1, v1 - v0)
        self.mode = "views"      # build() switches to "tracks" when it replicates the 16-byte records
        self.full = None         # "tracks": per level, the search structure over ALL views' records

    @classmethod
    def build(cls, fmaps_local, depths_local, intrs_local, extrs_local, v0: int, V_total: int, stride=4,
              levels: Sequence[int] = (0, 1, 2, 3), k=16, group=None, backend=None, reuse: Optional["ShardedWindow"] = None,
              intrs_all=None, extrs_all=None, mode: Optional[str] = None):
        """*_local: this rank's views only ([Vloc,S,C,H,W] ...); v0 = global index of the first one.
        mode "tracks" (default when intrs_all / extrs_all [V,S,3,3] / [V,S,3,4] are given): the 16-byte {x,y,z,depth}
        records of all views are all-gathered (3% of the cloud's bytes; features stay put) and every rank searches ALL
        views for its own track shard.  mode "views": nothing is replicated; every rank searches its own views for all
        tracks and candidate lists are merged (two bounded phases, see corr)."""
        if backend is None:
            from . import ops as backend
        if mode is None:
            mode = "tracks" if intrs_all is not None else "views"
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        if mode in ("p2p", "fused") and (world == 1 or V_total < world or not symmetric_memory_available()):
            mode = "tracks"         # needs >= 2 ranks, a view on every rank and symmetric memory
        Vloc = fmaps_local.shape[0]
        levels = list(levels)
        clouds = []
        if mode in ("p2p", "fused"):
            win = cls._build_p2p(fmaps_local, depths_local, intrs_local, extrs_local, v0, V_total, stride, levels, k, group,
                                 backend, reuse, intrs_all, extrs_all)
            win.mode = mode
            return win
        if Vloc > 0:
            clouds = backend.cloud_build(fmaps_local, depths_local, intrs_local, extrs_local, stride=stride, levels=levels,
                                         want_spacing=False, reuse=None if reuse is None else reuse.clouds)
        win = cls(clouds, v0, v0 + Vloc, V_total, k, group, backend, n_levels=len(levels))
        if reuse is not None:
            win._bufs = reuse._bufs
        win.mode = mode
        if mode == "tracks":
            win._replicate_records(fmaps_local, intrs_all, extrs_all, stride, levels)
        elif mode != "views":
            raise ValueError(mode)
        return win

    @classmethod
    def _build_p2p(cls, fmaps_local, depths_local, intrs_local, extrs_local, v0, V_total, stride, levels, k, group, backend, reuse,
                   intrs_all, extrs_all):
        """mode "p2p": cloud_build stores the 16-byte records of its views straight into every rank's replicated cloud
        (symmetric memory, multicast stores) -- no all-gather, no copy."""
        be = backend
        Vloc, S, H, W = fmaps_local.shape[0], fmaps_local.shape[1], fmaps_local.shape[3], fmaps_local.shape[4]
        dev = fmaps_local.device
        win = cls([], v0, v0 + Vloc, V_total, k, group, be, n_levels=len(levels))
        win.mode = "p2p"
        # whole 16-byte records cross NVLink by default; MVT_SPLIT_RECORDS=1: only the correlation values do and every rank
        # fills in the offsets from its replica (a quarter of the bytes -- measured equal at 8 GPUs, 3% slower at 2: the
        # exchange is bound by the number of small stores, not by their size)
        win.split_records = bool(os.environ.get("MVT_SPLIT_RECORDS"))
        if reuse is not None and reuse.mode in ("p2p", "fused"):
            win._bufs = reuse._bufs
        shared = {}
        for l in levels:
            hw = (H >> l) * (W >> l)
            shared[l] = win._shared(f"full{l}", (S, V_total * hw, 4), torch.float32, dev)
        first = win._shared("full%d" % levels[0], None, None, None)
        first.barrier()           # every rank is done reading the previous window's records
        win.clouds = be.cloud_build(fmaps_local, depths_local, intrs_local, extrs_local, stride=stride, levels=levels,
                                    want_spacing=False, reuse=None if (reuse is None or not reuse.clouds) else reuse.clouds,
                                    shared_records={l: shared[l].peers for l in levels}, view_offset=v0, V_total=V_total,
                                    want_tiles=False)
        first.barrier()           # ... and every rank's records have arrived everywhere
        cam_all = be.camera_prep(intrs_all, extrs_all)
        win.full = [be.full_cloud(l, shared[l].t, cam_all, V_total, H >> l, W >> l, float(stride) * (1 << l)) for l in levels]
        win._tiles()
        return win

    def _shared(self, name, shape, dtype, device):
        key = "shared:" + name
        b = self._bufs.get(key)
        if shape is None:
            return b
        if b is None or tuple(b.t.shape) != tuple(shape) or b.t.dtype != dtype:
            b = SharedBuffer(shape, dtype, device, self.group)      # collective: every rank allocates the same buffers
            self._bufs[key] = b
        return b

    def _replicate_records(self, fmaps_local, intrs_all, extrs_all, stride, levels):
        """All-gather the raster records of every level -> self.full[l]: search structure over all V views."""
        be, N, V = self.be, self.world, self.V_total
        S, H, W = fmaps_local.shape[1], fmaps_local.shape[3], fmaps_local.shape[4]
        dev = fmaps_local.device
        spans = [view_range(r, N, V) for r in range(N)]
        vmax = max(b - a for a, b in spans)
        even = all(b - a == vmax for a, b in spans)
        cam_all = be.camera_prep(intrs_all, extrs_all)
        self.full = []
        for i, l in enumerate(levels):
            Hl, Wl = H >> l, W >> l
            hw = Hl * Wl
            mine = self.clouds[i].xyz4 if self.clouds else None             # [S, Vloc*hw, 4]
            if N == 1:
                full = mine
            else:
                send = mine
                if not even:                                                  # pad to the largest shard
                    send = self._buf(f"send{l}", (S, vmax * hw, 4), torch.float32, dev)
                    if mine is not None:
                        send[:, :mine.shape[1]].copy_(mine)
                gathered = self._buf(f"gath{l}", (N, S, vmax * hw, 4), torch.float32, dev)
                dist.all_gather_into_tensor(gathered.view(N, -1), send.reshape(1, -1), group=self.group)
                full = self._buf(f"full{l}", (S, V * hw, 4), torch.float32, dev)
                if even:
                    full.view(S, N, vmax * hw, 4).copy_(gathered.permute(1, 0, 2, 3))
                else:
                    for r, (a, b) in enumerate(spans):
                        if b > a:
                            full[:, a * hw:b * hw].copy_(gathered[r, :, :(b - a) * hw])
            self.full.append(be.full_cloud(l, full, cam_all, V, Hl, Wl, float(stride) * (1 << l)))
        self._tiles()

    def _tiles(self):
        """Per-tile depth ranges of the replicated clouds (every rank builds its own: one small kernel over the records)."""
        build = getattr(self.be, "tile_depth_build", None)
        if build is None:
            return
        for c in self.full:
            shape = (c.xyz4.shape[0], c.V, (c.H + 7) // 8, (c.W + 7) // 8, 2)
            c.tile_z = self._buf(f"tiles{c.level}", shape, torch.float32, c.xyz4.device)
        build(self.full)

    def _buf(self, name, shape, dtype, device):
        t = self._bufs.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype or t.device != device:
            t = torch.empty(shape, dtype=dtype, device=device)
            self._bufs[name] = t
        return t

    def corr(self, targets: torch.Tensor, coords: torch.Tensor, out: Optional[torch.Tensor] = None, marks=None) -> torch.Tensor:
        """targets [S,M,C], coords [S,M,3] (replicated on every rank) -> [S, M, levels*k*4], identical on every rank.
        marks: optional list that receives a CUDA event after each of the six stages (profiling)."""
        def mark():
            if marks is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                marks.append(e)
        be, N, k, L = self.be, self.world, self.k, self.levels
        S, M, _ = targets.shape
        Ms = track_shard(N, M)
        dev = targets.device
        if self.mode == "tracks":
            return self._corr_tracks(targets, coords, out, mark)
        if self.mode == "p2p":
            return self._corr_p2p(targets, coords, out, mark)
        if self.mode == "fused":
            return self._corr_fused(targets, coords, out, mark)
        keys = self._buf("keys", (N, L, S, Ms, k), torch.int64, dev)
        recv = self._buf("recv", (N, L, S, Ms, k), torch.int64, dev) if N > 1 else keys
        offs = [self.v0 * c.H * c.W for c in self.clouds]
        caps = None
        if N > 1:
            # phase A: bounded-effort local searches; their merged k-th best bounds the GLOBAL k-th distance
            exact = self._buf("exact", (N, L, S, Ms), torch.uint8, dev)
            if self.clouds:
                be.knn_keys(self.clouds, offs, coords, k, n_shards=N, out=keys, budget=self.budget, exact_out=exact)
            else:
                keys.fill_(be.EMPTY_KEY)
                exact.fill_(1)
            _all_to_all(recv, keys, group=self.group)
            bound_mine = self._buf("bound_mine", (L * S * Ms, k), torch.int64, dev)
            be.topk_merge(recv.view(N, L * S * Ms, k), keys_out=bound_mine)
            caps = self._buf("caps", (N, L * S * Ms), torch.int64, dev)
            dist.all_gather_into_tensor(caps, bound_mine[:, k - 1].contiguous().view(1, -1), group=self.group)
        mark()
        # phase B: the local searches that gave up in phase A, exactly, within the bound (same key buffer)
        if self.clouds:
            be.knn_keys(self.clouds, offs, coords, k, n_shards=N, out=keys, caps=caps, skip=exact if N > 1 else None)
        else:
            keys.fill_(be.EMPTY_KEY)
        mark()
        if N > 1:
            _all_to_all(recv, keys, group=self.group)
        mark()
        idx_mine = self._buf("idx_mine", (L * S * Ms, k), torch.int32, dev)
        be.topk_merge(recv.view(N, L * S * Ms, k), out=idx_mine)
        mark()
        idx_all = self._buf("idx_all", (N, L, S, Ms, k), torch.int32, dev)
        if N > 1:
            dist.all_gather_into_tensor(idx_all.view(N, -1), idx_mine.view(1, -1), group=self.group)
        else:
            idx_all.view(-1).copy_(idx_mine.view(-1))
        mark()
        if out is None:
            out = torch.empty(S, M, L * k * 4, device=dev, dtype=torch.float32)
        if self.clouds:
            owned = [(c.xyz4, c.fvec, self.v0 * c.H * c.W, self.v1 * c.H * c.W) for c in self.clouds]
            be.corr_from_idx(owned, idx_all, targets, coords, k, out=out)
        else:
            out.zero_()
        mark()
        if N > 1:
            dist.all_reduce(out, op=dist.ReduceOp.SUM, group=self.group)
        mark()
        return out

    def _corr_tracks(self, targets, coords, out, mark):
        """mode "tracks": exact search of this rank's track shard over all views -> all-gather of the index lists ->
        correlation of the neighbours owned here -> all-reduce."""
        be, N, k, L, r = self.be, self.world, self.k, self.levels, self.rank
        S, M, _ = targets.shape
        Ms = track_shard(N, M)
        dev = targets.device
        lo, hi = r * Ms, min((r + 1) * Ms, M)
        mine = self._buf("coords_mine", (S, Ms, 3), torch.float32, dev)
        if hi > lo:
            mine[:, :hi - lo].copy_(coords[:, lo:hi])
        if hi - lo < Ms:                                   # padding tracks of the last shard(s): any valid position
            mine[:, max(hi - lo, 0):].copy_(coords[:, :1])
        keys = self._buf("keys_mine", (1, L, S, Ms, k), torch.int64, dev)
        be.knn_keys(self.full, [0] * L, mine, k, n_shards=1, out=keys)
        mark()
        idx_mine = self._buf("idx_mine", (L * S * Ms, k), torch.int32, dev)
        be.topk_merge(keys.view(1, L * S * Ms, k), out=idx_mine)
        mark()
        idx_all = self._buf("idx_all", (N, L, S, Ms, k), torch.int32, dev)
        if N > 1:
            dist.all_gather_into_tensor(idx_all.view(N, -1), idx_mine.view(1, -1), group=self.group)
        else:
            idx_all.view(-1).copy_(idx_mine.view(-1))
        mark()
        if out is None:
            out = torch.empty(S, M, L * k * 4, device=dev, dtype=torch.float32)
        if self.clouds:
            owned = [(c.xyz4, c.fvec, self.v0 * c.H * c.W, self.v1 * c.H * c.W) for c in self.clouds]
            be.corr_from_idx(owned, idx_all, targets, coords, k, out=out)
        else:
            out.zero_()
        mark()
        if N > 1:
            dist.all_reduce(out, op=dist.ReduceOp.SUM, group=self.group)
        mark()
        return out

    def _corr_p2p(self, targets, coords, out, mark):
        """mode "p2p": the exchanges are stores into peer memory issued by the producing kernels (multicast when the
        switch offers it); two ~6 us cross-rank barriers per iteration replace the NCCL collectives."""
        be, N, k, L, r = self.be, self.world, self.k, self.levels, self.rank
        S, M, _ = targets.shape
        Ms = track_shard(N, M)
        dev = targets.device
        lo, hi = r * Ms, min((r + 1) * Ms, M)
        mine = self._buf("coords_mine", (S, Ms, 3), torch.float32, dev)
        if hi > lo:
            mine[:, :hi - lo].copy_(coords[:, lo:hi])
        if hi - lo < Ms:
            mine[:, max(hi - lo, 0):].copy_(coords[:, :1])
        idx_all = self._shared("idx_all", (N, L, S, Ms, k), torch.int32, dev)
        flip = 1 - self._bufs.get("flip", 1)       # two result buffers alternate: a rank may still be reading the last one
        self._bufs["flip"] = flip
        res = self._shared(f"out{flip}", (S, M, L * k * 4), torch.float32, dev)
        be.knn_idx_shared(self.full, mine, k, r, N, idx_all.peers)
        mark()
        idx_all.barrier()
        mark()
        owned = [(c.xyz4, c.fvec, self.v0 * c.H * c.W, self.v1 * c.H * c.W) for c in self.clouds]
        be.corr_from_idx_shared(owned, idx_all.t, targets, coords, k, r, res.peers, res.t.stride(1),
                                replicas=[c.xyz4 for c in self.full] if self.split_records else None)
        mark()
        res.barrier()
        mark()
        if out is not None:
            out.copy_(res.t)
            return out
        return res.t

    def _corr_fused(self, targets, coords, out, mark):
        """mode "fused": ONE persistent kernel per rank and iteration (mvt_shard_iter): it searches this rank's track shard,
        pushes one 8-byte record per neighbour into the inbox of the rank that owns the neighbour's features, correlates the
        records arriving in its own inbox and stores the results into every rank's block; the kernel ends with a cross-rank
        epoch exchange, so there is no barrier launch and no list walk (mvtracker_b200/csrc/mvt_shard.cu)."""
        be, N, k, L = self.be, self.world, self.k, self.levels
        S, M, _ = targets.shape
        dev = targets.device
        st = self._bufs.get("shard_state")
        Ms = track_shard(N, M)
        cap = L * S * Ms * k
        if st is None or st.cap != cap or st.results[0][0].shape != (S, M, L * k * 4):
            inbox = self._shared("inbox", (N, cap), torch.int64, dev)
            ctrl = self._shared("ctrl", (16,), torch.int64, dev)
            res = [self._shared(f"out{i}", (S, M, L * k * 4), torch.float32, dev) for i in range(2)]
            inbox.t.zero_(); ctrl.t.zero_()
            torch.cuda.current_stream().synchronize()
            inbox.barrier()
            spans = [view_range(r, N, self.V_total) for r in range(N)]
            st = be.ShardState(self.rank, N, [a for a, _ in spans] + [self.V_total], (inbox.t, inbox.peers), (ctrl.t, ctrl.peers),
                               [(r.t, r.peers) for r in res], cap, dev)
            self._bufs["shard_state"] = st
        res = be.shard_iter(self.full, self.clouds, targets, coords, k, st)
        mark()
        if out is not None:
            out.copy_(res)
            return out
        return res


def _offset_peers(peers, nbytes: int):
    """The same symmetric buffer, every mapping (and the multicast address) advanced by nbytes."""
    from ._lib import Peers
    q = Peers()
    q.n = peers.n
    q.mc = (peers.mc + nbytes) if peers.mc else None
    for r in range(peers.n):
        q.ptr[r] = peers.ptr[r] + nbytes
    return q


class ShardedSequence:
    """Sliding windows of ONE sequence across ranks with the per-frame cloud cache (the multi-GPU WindowSequence): rank r
    keeps the clouds of ITS views in a local CloudArena and every rank keeps the replicated 16-byte search records of ALL
    views in a symmetric-memory arena of the same slots.  push() builds only the frames a window adds -- the producing
    kernel stores their records into every rank's arena (multicast) -- and returns a ShardedWindow over the current run."""

    def __init__(self, V_total: int, v0: int, v1: int, H: int, W: int, C: int, dtype, device, S: int = 12, hop: Optional[int] = None,
                 stride=4, levels: Sequence[int] = (0, 1, 2, 3), k=16, slots: Optional[int] = None, group=None, backend=None,
                 mode: str = "fused"):
        if backend is None:
            from . import ops as backend
        self.be, self.group, self.mode = backend, group, mode
        self.V, self.v0, self.v1, self.H, self.W = V_total, v0, v1, H, W
        self.S, self.hop, self.stride, self.levels, self.k = int(S), int(hop or S // 2), float(stride), list(levels), k
        if slots is None:
            per = backend.CloudArena.bytes_per_frame(v1 - v0, H, W, C, torch.empty(0, dtype=dtype).element_size(), levels) \
                + sum(V_total * (H >> l) * (W >> l) * 16 for l in levels)
            slots = int(max(self.S + self.hop, min(4 * self.S, (torch.cuda.mem_get_info(device)[0] // 3) // max(per, 1))))
            t = torch.tensor([slots], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)      # the symmetric arenas must have one size
            slots = int(t.item())
        self.slots = slots
        self.arena = backend.CloudArena(v1 - v0, H, W, C, dtype, device, levels, slots, stride, want_valid=False)
        self.rec = {l: SharedBuffer((slots, V_total * (H >> l) * (W >> l), 4), torch.float32, device, group) for l in levels}
        self.tiles = {l: torch.empty(slots, V_total, ((H >> l) + 7) // 8, ((W >> l) + 7) // 8, 2, device=device) for l in levels}
        self.win = ShardedWindow([], v0, v1, V_total, k, group, backend, n_levels=len(self.levels))
        self.win.mode = mode
        self.slot0 = None
        self.frames_built = 0

    def push(self, fmaps_new, depths_new, intrs_local, extrs_local, intrs_all, extrs_all) -> "ShardedWindow":
        """fmaps_new [Vloc, n, C, H, W] / depths_new [Vloc, n, H, W]: the frames this window adds, THIS rank's views (all S of the
        first window, `hop` afterwards); intrs / extrs _local [Vloc, S, ...] and _all [V, S, ...]: cameras of the whole window."""
        be, S, hop, ar = self.be, self.S, self.hop, self.arena
        n = fmaps_new.shape[1]
        any_rec = self.rec[self.levels[0]]
        any_rec.barrier()                       # every rank has finished reading the window that is about to change
        if self.slot0 is None:
            assert n == S, f"the first window needs all {S} frames"
            self.slot0, first_new = 0, 0
        else:
            assert n == hop, f"a later window adds {hop} frames"
            self.slot0 += hop
            if self.slot0 + S > self.slots:
                ar.move(self.slot0, 0, S - hop)
                for l in self.levels:           # (each rank moves its own replica; nobody writes it before the barrier below)
                    self.rec[l].t[0:S - hop].copy_(self.rec[l].t[self.slot0:self.slot0 + S - hop])
                    self.tiles[l][0:S - hop].copy_(self.tiles[l][self.slot0:self.slot0 + S - hop])
                self.slot0 = 0
                any_rec.barrier()
            first_new = S - hop
        s_new = self.slot0 + first_new
        cam_loc = be.camera_prep(intrs_local, extrs_local)
        cam_run = cam_loc.reshape(-1)[first_new * cam_loc.shape[1]:]
        tgt = ar.run(s_new, n, cam_run, tiles_ready=False)
        shared = {l: _offset_peers(self.rec[l].peers, s_new * self.rec[l].t[0].numel() * 4) for l in self.levels}
        be.cloud_build(fmaps_new, depths_new, None, None, stride=self.stride, levels=self.levels, want_spacing=False, cam=cam_run,
                       reuse=tgt, shared_records=shared, view_offset=self.v0, V_total=self.V, cam_view_stride=S)
        any_rec.barrier()                       # every rank's records of the new frames have arrived everywhere
        cam_all = be.camera_prep(intrs_all, extrs_all)
        # tile tables of the new frames of the replicated records (every rank, its own replica)
        new_full = [be.Cloud(l, None, None, None, None, self.rec[l].t[s_new:s_new + n], cam_all.reshape(-1)[first_new * cam_all.shape[1]:],
                             self.V, self.H >> l, self.W >> l, self.stride * (1 << l), self.tiles[l][s_new:s_new + n]) for l in self.levels]
        be.tile_depth_build(new_full, cam_view_stride=S)
        sl = slice(self.slot0, self.slot0 + S)
        self.win.clouds = ar.run(self.slot0, S, cam_loc)
        self.win.full = [be.Cloud(l, None, None, None, None, self.rec[l].t[sl], cam_all, self.V, self.H >> l, self.W >> l,
                                  self.stride * (1 << l), self.tiles[l][sl], True) for l in self.levels]
        self.frames_built += n
        return self.win


def frame_owner(t: int, world: int) -> int:
    """Rank that owns global frame t: round robin, so that the 12 consecutive frames of a window AND the 6 frames an advance
    adds spread over the ranks, and a frame keeps its owner (and its cached cloud) while windows slide over it."""
    return t % world


class FrameShardedSequence:
    """Sliding windows of ONE sequence with the FRAMES of a window sharded over the ranks (mode "frames").

    The frames of a window are the batch axis of the reference's kNN correlation (mvtracker.py:800-846 runs on [B*S, ...]:
    frame s of the tracks only meets frame s of the cloud): independent units.  Rank r owns the global frames t with
    t % world == r, builds and caches THEIR clouds of all views (a local CloudArena: nothing of a cloud ever crosses NVLink,
    neither features nor search records), and per refinement iteration runs the single-GPU fused kernel on its frames x all
    tracks; the kernel's epilogue stores every 16-byte result record into the [S, M, levels*k*4] block of EVERY rank
    (multimem.st through the NVSwitch multicast address), one signal-pad barrier completes the block.  Per iteration the
    node moves S*M*levels*k*16 bytes to each rank (50 MB for the panoptic window) and nothing else.

        seq = FrameShardedSequence(V, H', W', C, dtype, device)
        ts = seq.new_frames()                                  # global frame numbers this rank has to supply next
        win = seq.push(fmaps[:, ts], depths[:, ts], intrs_win, extrs_win)
        fcorrs = win.corr(ffeats, coords)                      # [S, M, width], complete on every rank

    Results are bit-identical to the single-GPU kernel (it IS that kernel, on a subset of the frames)."""

    def __init__(self, V: int, H: int, W: int, C: int, dtype, device, S: int = 12, hop: Optional[int] = None, stride=4,
                 levels: Sequence[int] = (0, 1, 2, 3), k=16, slots: Optional[int] = None, group=None, backend=None,
                 rank: Optional[int] = None, world: Optional[int] = None, shared_out: Optional[torch.Tensor] = None):
        if backend is None:
            from . import ops as backend
        self.be, self.group = backend, group
        live = dist.is_initialized() and world is None
        self.world = world if world is not None else (dist.get_world_size(group) if live else 1)
        self.rank = rank if rank is not None else (dist.get_rank(group) if live else 0)
        self.S, self.hop, self.k, self.levels = int(S), int(hop or S // 2), k, list(levels)
        self.width = len(self.levels) * k * 4
        most = (self.S + self.world - 1) // self.world              # frames of one window a rank can own
        if slots is None:
            slots = 4 * most
        if slots < 2 * most:
            raise ValueError(f"{most} frames per window and rank need at least {2 * most} slots")
        self.arena = backend.CloudArena(V, H, W, C, dtype, device, self.levels, slots, stride, want_valid=False)
        self.device = device
        self.shared_out = shared_out        # emulated ranks of one process: all of them store into this block, no exchange
        if shared_out is not None or self.world == 1:
            self.transport = "local"
        elif symmetric_memory_available():
            self.transport = "multicast"
        else:
            self.transport = "allreduce"    # gloo / no symmetric memory: zeros + the local frames' records, summed
        self.mode = "frames"
        self.windows, self.frames_built = 0, 0
        self.slot0, self.run = 0, []        # global frame numbers of the current run, slot of run[i] = slot0 + i
        self.clouds, self.plan, self.pos0 = [], None, 0
        self._res, self._flip = {}, 0
        self._pending, self._side, self.last_build = None, None, None

    @property
    def has_prefetch(self) -> bool:
        """A prefetch() is waiting for its push()."""
        return self._pending is not None

    # ---- which frames --------------------------------------------------------------------------------------------------------
    def window_frames(self, w: Optional[int] = None) -> range:
        w = self.windows if w is None else w
        return range(self.hop * w, self.hop * w + self.S)

    def new_frames(self) -> List[int]:
        """Global numbers of the frames the next push() adds AND this rank owns (ascending; may be empty)."""
        w = self.windows
        added = self.window_frames(w) if w == 0 else range(self.hop * w + self.S - self.hop, self.hop * w + self.S)
        return [t for t in added if frame_owner(t, self.world) == self.rank]

    # ---- build ---------------------------------------------------------------------------------------------------------------
    def _stage(self, fmaps_new, depths_new, intrs_win, extrs_win):
        """Builds (on the current stream) what the next window adds and returns that window's state; self is not changed."""
        be, ar = self.be, self.arena
        new = self.new_frames()
        n = 0 if fmaps_new is None else fmaps_new.shape[1]
        if n != len(new):
            raise ValueError(f"rank {self.rank} adds frames {new} to window {self.windows}, got {n} frames")
        t0 = self.hop * self.windows
        kept = [t for t in self.run if t >= t0]                    # a suffix of the run: frames the windows share
        slot0 = self.slot0 + len(self.run) - len(kept)
        if slot0 + len(kept) + n > ar.slots:                        # wrap: the kept frames go to the front
            if kept:
                ar.move(slot0, 0, len(kept))
            slot0 = 0
        run = kept + new
        nw = len(run)
        pos0 = (run[0] - t0) if nw else 0                           # window positions pos0, pos0 + world, ...
        clouds, plan = [], None
        if nw:
            # the local frames' camera records, record v * nw + i (a strided slice of the window's: no index tensor, no host copy)
            cam_win = be.camera_prep(intrs_win, extrs_win)
            V = intrs_win.shape[0]
            cam = cam_win.view(V, self.S, -1)[:, pos0::self.world][:, :nw].contiguous().view(V * nw, -1)
            if n:
                first_new = len(kept)
                ar.build(slot0 + first_new, fmaps_new, depths_new, cam.reshape(-1)[first_new * cam.shape[1]:], nw)
            clouds = ar.run(slot0, nw, cam)
            make = getattr(be, "CorrPlan", None)       # the kernel's per-level descriptors, made once per window
            lv = [(c, c.fvec, self.k) for c in clouds]
            plan = make(lv) if make is not None else lv
        return dict(slot0=slot0, run=run, pos0=pos0, clouds=clouds, plan=plan, n=n)

    def _adopt(self, st) -> "FrameShardedSequence":
        self.slot0, self.run, self.pos0, self.clouds, self.plan = st["slot0"], st["run"], st["pos0"], st["clouds"], st["plan"]
        self.windows += 1
        self.frames_built += st["n"]
        return self

    def push(self, fmaps_new=None, depths_new=None, intrs_win=None, extrs_win=None) -> "FrameShardedSequence":
        """fmaps_new [V, n, C, H', W'] / depths_new [V, n, H', W'] (or [V, n, 1, H', W']): the frames new_frames() named, ALL
        views (None / n = 0 when this rank owns none of the added frames); intrs_win [V, S, 3, 3] / extrs_win [V, S, 3, 4]:
        the cameras of the whole window (replicated: a few kilobytes).  After prefetch(...): push() without arguments."""
        if self._pending is not None:
            st, done = self._pending
            self._pending = None
            if st is None:                      # prefetch() could not run ahead: `done` holds its arguments, build now
                return self._adopt(self._stage(*done))
            torch.cuda.current_stream().wait_event(done)
            return self._adopt(st)
        if intrs_win is None:
            raise ValueError("push() without arguments needs a prefetch() before it")
        return self._adopt(self._stage(fmaps_new, depths_new, intrs_win, extrs_win))

    def prefetch(self, fmaps_new, depths_new, intrs_win, extrs_win) -> None:
        """push()'s work for the NEXT window, started now on a side stream: its frames are built (and, at a wrap-around, the
        frames it keeps are moved) while the current window's iterations run -- those only read the current run, and a rank's
        correlation kernels leave the GPU idle at every barrier and in their tails (at N = 8 a rank computes 1-2 frames per
        iteration).  The inputs must stay untouched until the push() that follows.  Needs an arena of >= 3 windows' worth of
        local frames (the next run must not overlap the current one); a smaller arena builds at push() as before."""
        most = (self.S + self.world - 1) // self.world
        if self._pending is not None:
            raise RuntimeError("one prefetch() per push()")
        if self.device is None or torch.device(self.device).type != "cuda" or self.arena.slots < 3 * most or self.windows == 0:
            self._pending = (None, (fmaps_new, depths_new, intrs_win, extrs_win))       # built by the push() that follows
            return
        if self._side is None:
            self._side = torch.cuda.Stream(device=self.device)
        main = torch.cuda.current_stream()
        self._side.wait_stream(main)                # inputs are ready; earlier readers of the slots about to be written are done
        with torch.cuda.stream(self._side):
            t_a = torch.cuda.Event(enable_timing=True); t_a.record()
            st = self._stage(fmaps_new, depths_new, intrs_win, extrs_win)
            done = torch.cuda.Event(enable_timing=True); done.record()
        for c in st["clouds"]:
            if c.cam is not None:
                c.cam.record_stream(main)           # allocated on the side stream, read by the main stream's kernels later
        self.last_build = (t_a, done)
        self._pending = (st, done)

    # ---- one refinement iteration ----------------------------------------------------------------------------------------------
    def _result(self, S: int, M: int):
        key = (S, M)
        if key not in self._res:
            if self.transport == "multicast":
                self._res[key] = [SharedBuffer((S, M, self.width), torch.float32, self.device, self.group) for _ in range(2)]
            else:
                self._res[key] = [torch.empty(S, M, self.width, device=self.device, dtype=torch.float32) for _ in range(2)]
        self._flip ^= 1
        return self._res[key][self._flip]

    def corr(self, targets: torch.Tensor, coords: torch.Tensor, out: Optional[torch.Tensor] = None, marks=None) -> torch.Tensor:
        """targets [S, M, C], coords [S, M, 3] (all frames, replicated) -> [S, M, levels*k*4], complete on every rank.
        Two result blocks alternate: the block an iteration stores into was last read two iterations ago, and every rank has
        passed the barrier of the iteration in between, so one barrier per iteration orders everything."""
        be = self.be
        S, M, _ = targets.shape
        lv = self.plan
        if self.transport == "local":
            res = self.shared_out if self.shared_out is not None else self._result(S, M)
            if self.clouds:
                be.knn_corr_shared(lv, targets, coords, self.pos0, self.world, be.plain_peers(res), self.width)
        elif self.transport == "multicast":
            buf = self._result(S, M)
            if self.clouds:
                be.knn_corr_shared(lv, targets, coords, self.pos0, self.world, buf.peers, self.width)
            if marks is not None:
                marks.append(torch.cuda.Event(enable_timing=True)); marks[-1].record()
            buf.barrier()
            res = buf.t
        else:
            res = self._result(S, M)
            res.zero_()
            if self.clouds:
                be.knn_corr_shared(lv, targets, coords, self.pos0, self.world, be.plain_peers(res), self.width)
            dist.all_reduce(res, group=self.group)      # one non-zero contributor per record: the sum is exact
        if out is not None:
            out.copy_(res)
            return out
        return res
