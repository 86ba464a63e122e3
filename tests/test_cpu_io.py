"""CPU: the data formats either side of the path: demo sample / result .npz schema (demo.py:650-670,
1093-1118), DROID depth and extrinsics conventions (generate_npz_from_external_cams.py:231-250), window schedule."""
import numpy as np
import pytest
import torch

from mvtracker_b200 import io as mio


def _sample(V=2, T=5, H=6, W=8, N=3, seed=0):
    r = np.random.default_rng(seed)
    return dict(rgbs=r.integers(0, 255, (V, T, 3, H, W)).astype(np.uint8), depths=r.random((V, T, 1, H, W)).astype(np.float32),
                intrs=np.tile(np.eye(3, dtype=np.float32), (V, T, 1, 1)), extrs=np.tile(np.eye(4, dtype=np.float32)[:3], (V, T, 1, 1)),
                query_points=r.random((N, 4)).astype(np.float32), camera_ids=np.array(["cam_a", "cam_b"]))


def test_sample_roundtrip_and_result_schema(tmp_path):
    s = _sample()
    p = str(tmp_path / "sample.npz")
    np.savez_compressed(p, **s)
    got = mio.load_sample_npz(p)
    assert got["rgbs"].dtype == torch.float32 and tuple(got["rgbs"].shape) == (2, 5, 3, 6, 8)
    assert torch.equal(got["depths"], torch.from_numpy(s["depths"])) and got["camera_ids"] == ["cam_a", "cam_b"]
    T, N = 5, 3
    out = str(tmp_path / "result.npz")
    mio.save_tracks_npz(out, torch.zeros(T, N, 3), torch.ones(T, N), got["query_points"], got)
    z = np.load(out, allow_pickle=True)
    for k in ("tracks_3d", "visibilities", "query_points", "rgbs", "depths", "intrs", "extrs", "camera_ids", "tracker",
              "temporal_stride", "spatial_downsample"):            # the keys demo.py:1093-1118 writes
        assert k in z.files, k
    assert z["tracks_3d"].shape == (T, N, 3) and z["visibilities"].shape == (T, N) and z["extrs"].shape == (2, 5, 3, 4)


@pytest.mark.parametrize("bad", ["no_depths", "depth_shape", "extrs_4x4", "query_cols"])
def test_sample_validation(bad):
    s = _sample()
    if bad == "no_depths":
        del s["depths"]
    elif bad == "depth_shape":
        s["depths"] = s["depths"][:, :, 0]
    elif bad == "extrs_4x4":
        s["extrs"] = np.tile(np.eye(4, dtype=np.float32), (2, 5, 1, 1))
    else:
        s["query_points"] = s["query_points"][:, :3]
    with pytest.raises(mio.SampleFormatError):
        mio.check_sample(s)


def test_droid_conventions():
    r = np.random.default_rng(1)
    d = r.random((2, 3, 4, 5)).astype(np.float32)
    d[0, 0, 0, 0], d[1, 2, 3, 4], d[0, 1, 2, 3] = np.nan, np.inf, -np.inf
    c2w = np.tile(np.eye(4), (2, 3, 1, 1))
    c2w[..., :3, 3] = r.random((2, 3, 3))
    ang = 0.3
    c2w[0, :, :3, :3] = np.array([[np.cos(ang), -np.sin(ang), 0], [np.sin(ang), np.cos(ang), 0], [0, 0, 1]])
    depths, w2c = mio.droid_to_reference(d, c2w)
    assert depths.shape == (2, 3, 1, 4, 5) and np.isfinite(depths).all()
    assert depths[0, 0, 0, 0, 0] == 0 and depths[1, 2, 0, 3, 4] == 0 and depths[0, 1, 0, 2, 3] == 0
    assert w2c.shape == (2, 3, 3, 4)
    full = np.concatenate([w2c, np.tile(np.array([0, 0, 0, 1.0]), (2, 3, 1, 1))], axis=-2)
    assert np.allclose(full @ c2w, np.eye(4), atol=1e-12)
    c2w[0, 0, 0, 0] = np.nan
    with pytest.raises(mio.SampleFormatError):
        mio.droid_to_reference(d, c2w)


def test_window_schedule_matches_reference_counts():
    # 3 / 10 / 16 / 24 windows for T = 24 / 64 / 100 / 150 (while w < T - S/2, step S/2; mvtracker.py:522)
    assert [len(mio.window_starts(T)) for T in (24, 64, 100, 150)] == [3, 10, 16, 24]
    assert mio.window_starts(24) == [0, 6, 12] and mio.window_starts(5) == [0]
