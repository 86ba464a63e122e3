"""Data formats either side of the path: the reference's demo `.npz` sample schema, its result
schema, the DROID depth / extrinsics conventions, and pinned-host streaming of a sequence's sliding windows.

Host-side only (numpy + torch tensors); nothing here computes on the hot path.
  sample in :  demo.py:650-670   rgbs [V,T,3,H,W], depths [V,T,1,H,W], intrs [V,T,3,3], extrs [V,T,3,4] (world -> camera),
               optional query_points [N,4] (t, x, y, z), camera_ids [V], timestamps / per_camera_timestamps
  result out:  demo.py:1093-1118 tracks_3d [T,N,3], visibilities [T,N], query_points [N,4] + the sample's camera data + metadata
  DROID:       conversions/droid/debug/generate_npz_from_external_cams.py:231 (NaN / +-Inf depth -> 0 = invalid),
               :243 (camera->world 4x4 inverted to world->camera), :250 (cropped to 3x4)
  windows:     mvtracker/models/core/mvtracker/mvtracker.py:522-566 (S = 12 frames, step S/2, last window padded by repeating
               the final frame)
"""
from typing import Dict, Iterator, Optional, Tuple

import numpy as np
import torch

SAMPLE_KEYS = ("rgbs", "depths", "intrs", "extrs")
OPTIONAL_KEYS = ("query_points", "camera_ids", "timestamps", "per_camera_timestamps")


class SampleFormatError(ValueError):
    pass


def check_sample(sample: Dict[str, np.ndarray]) -> Tuple[int, int, int, int]:
    """Validates the demo schema; -> (V, T, H, W)."""
    for k in SAMPLE_KEYS:
        if k not in sample:
            raise SampleFormatError(f"sample has no '{k}' (keys: {sorted(sample.keys())})")
    rgbs, depths, intrs, extrs = (sample[k] for k in SAMPLE_KEYS)
    if rgbs.ndim != 5 or rgbs.shape[2] != 3:
        raise SampleFormatError(f"rgbs must be [V,T,3,H,W], got {rgbs.shape}")
    V, T, _, H, W = rgbs.shape
    if tuple(depths.shape) != (V, T, 1, H, W):
        raise SampleFormatError(f"depths must be {(V, T, 1, H, W)}, got {tuple(depths.shape)}")
    if tuple(intrs.shape) != (V, T, 3, 3):
        raise SampleFormatError(f"intrs must be {(V, T, 3, 3)}, got {tuple(intrs.shape)}")
    if tuple(extrs.shape) != (V, T, 3, 4):
        raise SampleFormatError(f"extrs must be {(V, T, 3, 4)} (world -> camera), got {tuple(extrs.shape)}")
    if "query_points" in sample and (sample["query_points"].ndim != 2 or sample["query_points"].shape[1] != 4):
        raise SampleFormatError(f"query_points must be [N,4] (t,x,y,z), got {sample['query_points'].shape}")
    return V, T, H, W


def load_sample_npz(path: str, mmap: bool = True) -> Dict[str, object]:
    """Reads a demo sample (demo.py:650: np.load(..., mmap_mode='r', allow_pickle=True)) -> dict of float32 torch tensors on
    the host (camera_ids stays a list of str).  Depth that is NaN / Inf is left as stored; see droid_to_reference."""
    with np.load(path, mmap_mode="r" if mmap else None, allow_pickle=True) as z:
        raw = {k: z[k] for k in z.files}
    check_sample(raw)
    out: Dict[str, object] = {k: torch.from_numpy(np.ascontiguousarray(raw[k], dtype=np.float32)) for k in SAMPLE_KEYS}
    if "query_points" in raw:
        out["query_points"] = torch.from_numpy(np.ascontiguousarray(raw["query_points"], dtype=np.float32))
    if "camera_ids" in raw:
        out["camera_ids"] = [x.decode() if isinstance(x, bytes) else str(x) for x in np.asarray(raw["camera_ids"]).reshape(-1)]
    for k in ("timestamps", "per_camera_timestamps"):
        if k in raw:
            out[k] = np.asarray(raw[k])
    return out


def save_tracks_npz(path: str, tracks_3d, visibilities, query_points, sample: Dict[str, object], tracker: str = "mvtracker",
                    temporal_stride: int = 1, spatial_downsample: int = 1) -> None:
    """Writes the result file demo.py:1093-1118 writes (np.savez_compressed; the keys refine_tracks.py reads)."""
    to_np = lambda t: t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
    tr, vi, qp = to_np(tracks_3d), to_np(visibilities), to_np(query_points)
    if tr.ndim != 3 or tr.shape[2] != 3 or vi.shape != tr.shape[:2] or qp.shape != (tr.shape[1], 4):
        raise SampleFormatError(f"tracks_3d [T,N,3] / visibilities [T,N] / query_points [N,4] expected, got {tr.shape}, {vi.shape}, {qp.shape}")
    data = {"tracks_3d": tr, "visibilities": vi, "query_points": qp}
    for k in SAMPLE_KEYS:
        data[k] = to_np(sample[k])
    data["camera_ids"] = np.asarray(sample.get("camera_ids")) if sample.get("camera_ids") is not None else None
    data.update(tracker=tracker, temporal_stride=temporal_stride, spatial_downsample=spatial_downsample)
    for k in ("per_camera_timestamps", "timestamps"):
        if k in sample:
            data[k] = sample[k]
    np.savez_compressed(path, **data)


def droid_to_reference(depths: np.ndarray, cam_to_world: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """DROID conventions -> the model's: depths [V,T,H,W] or [V,T,1,H,W] with NaN / +-Inf -> 0 (0 = invalid depth, which
    cloud_build marks invalid); cam_to_world [V,T,4,4] -> world -> camera [V,T,3,4].  Raises on non-finite extrinsics, as the
    reference's converter does (generate_npz_from_external_cams.py:238-247)."""
    d = np.nan_to_num(np.asarray(depths, dtype=np.float32), nan=0.0, posinf=0.0, neginf=0.0)
    if d.ndim == 4:
        d = d[:, :, None]
    e = np.asarray(cam_to_world)
    if e.shape[-2:] != (4, 4):
        raise SampleFormatError(f"cam_to_world must be [...,4,4], got {e.shape}")
    if not np.isfinite(e).all():
        raise SampleFormatError("extrinsics contain NaN or Inf values")
    w2c = np.linalg.inv(e)
    if not np.isfinite(w2c).all():
        raise SampleFormatError("inverted extrinsics contain NaN or Inf values")
    return d, np.ascontiguousarray(w2c[..., :3, :])


def window_starts(T: int, S: int = 12) -> list:
    """First frames of the model's sliding windows: 0, S/2, ... while start < T - S/2 (mvtracker.py:522; 3 / 10 / 16 / 24
    windows for T = 24 / 64 / 100 / 150)."""
    step = S // 2
    return list(range(0, max(T - step, 1), step))


class WindowStream:
    """Streams a host-resident sequence window by window to the GPU: the frames of window w+1 are copied from pinned host
    memory on a side stream while window w is being processed (the reference slices device tensors that hold the whole
    sequence, mvtracker.py:559-566; a 150-frame 8-view fp32 feature sequence does not need to be resident).

    tensors: dict name -> [V, T, ...] host tensor (sliced along dim 1).  Yields (start, dict name -> [V, S, ...] device
    tensor); the last window is padded by repeating the final frame (mvtracker.py:561-566)."""

    def __init__(self, tensors: Dict[str, torch.Tensor], S: int = 12, device: str = "cuda:0"):
        self.S, self.device = S, torch.device(device)
        self.T = next(iter(tensors.values())).shape[1]
        self.host = {k: (t if t.is_pinned() else t.contiguous().pin_memory()) for k, t in tensors.items()}
        self.copy_stream = torch.cuda.Stream(device=self.device)

    def _issue(self, start: int):
        S, T = self.S, self.T
        out = {}
        with torch.cuda.stream(self.copy_stream):
            for k, t in self.host.items():
                n = min(S, T - start)
                dev = torch.empty((t.shape[0], S) + tuple(t.shape[2:]), dtype=t.dtype, device=self.device)
                dev[:, :n].copy_(t[:, start:start + n], non_blocking=True)
                if n < S:
                    dev[:, n:] = dev[:, n - 1:n]
                out[k] = dev
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        return out, ev

    def __iter__(self) -> Iterator[Tuple[int, Dict[str, torch.Tensor]]]:
        starts = window_starts(self.T, self.S)
        nxt = self._issue(starts[0])
        for i, st in enumerate(starts):
            cur, ev = nxt
            if i + 1 < len(starts):
                nxt = self._issue(starts[i + 1])
            torch.cuda.current_stream(self.device).wait_event(ev)
            for t in cur.values():
                t.record_stream(torch.cuda.current_stream(self.device))
            yield st, cur
