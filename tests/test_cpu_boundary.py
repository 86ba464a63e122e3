"""CPU-only checks of the boundary: the C-ABI library loads and exports every symbol include/mvt.h declares, argument
validation works without touching the GPU, the host mirror refuses CPU tensors (no fallback), and install() rebinds the
reference's seam names."""
import ctypes
import os
import re

import pytest
import torch

from mvtracker_b200 import _lib, dropin, ops
from oracle import ref_import

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "mvt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mvt_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = _lib.lib()
    names = _declared_symbols()
    assert len(names) >= 10
    for n in names:
        assert hasattr(L, n), f"libmvt.so does not export {n}"
    assert set(names) == set(_lib.SYMBOLS)
    assert L.mvt_version() == 100


def test_workspace_size_is_monotone_and_aligned():
    L = _lib.lib()
    prev = 0
    for n in (1, 100, 4096, 172800, 1 << 20):
        b = L.mvt_grid_workspace_bytes(12, n)
        assert b % 256 == 0 and b > prev
        prev = b
    assert L.mvt_grid_workspace_bytes(0, 10) == 0


def test_argument_errors_are_reported_without_a_gpu():
    L = _lib.lib()
    dummy = ctypes.c_void_p(256)
    assert L.mvt_knn(None, 1, None, 5, 16, None, 0, None, None) == -1
    assert b"null" in L.mvt_last_error()
    sr = _lib.Search(256, 256, None, None, 0, 10, 0, 0, 0, 0.0, 0)           # grid mode
    assert L.mvt_knn(ctypes.byref(sr), 1, dummy, 5, 64, None, 0, None, None) == -1      # k > 32
    assert b"k must be" in L.mvt_last_error()
    empty = _lib.Search(None, None, None, None, 0, 10, 0, 0, 0, 0.0, 0)
    assert L.mvt_knn(ctypes.byref(empty), 1, dummy, 5, 8, None, 0, None, None) == -1
    assert b"neither a grid nor raster" in L.mvt_last_error()
    bad = _lib.Search(None, None, 256, 256, 12, 100, 3, 5, 5, 4.0, 0)        # N != V*H*W
    assert L.mvt_knn(ctypes.byref(bad), 1, dummy, 5, 8, None, 0, None, None) == -1
    assert b"V*H*W" in L.mvt_last_error()
    lv = (_lib.CorrLevel * 4)()
    lv[0] = _lib.CorrLevel(_lib.Search(256, 256, None, None, 0, 8, 0, 0, 0, 0.0, 0), 256, 16, 0)
    rc = L.mvt_knn_corr(lv, 1, 1, 4, 128, 1, 0, dummy, dummy, 1, 0, dummy, 64, 0, None, None, None, 0, None)
    assert rc == -1 and b"k should be less than or equal to N" in L.mvt_last_error()   # mvtracker.py:785
    rc = L.mvt_knn_corr(lv, 1, 1, 4, 128, 3, 0, dummy, dummy, 1, 0, dummy, 64, 0, None, None, None, 0, None)
    assert rc == -1 and b"groups" in L.mvt_last_error()                                 # mvtracker.py:786-787
    outs = (_lib.LevelOut * 4)()
    assert L.mvt_cloud_build(dummy, 0, dummy, dummy, 2, 2, 8, 16, 16, 4.0, 0, outs, 0x10, None) == -1   # level 4


def test_no_cpu_fallback():
    x = torch.zeros(1, 10, 3)
    with pytest.raises(_lib.MvtError):
        dropin.knn(1, x, x)
    with pytest.raises(_lib.MvtError):
        dropin.PointcloudCorrBlock(1, 1, x, torch.zeros(1, 10, 4), True, False)
    with pytest.raises(_lib.MvtError):
        dropin.init_pointcloud_from_rgbd(torch.zeros(1, 1, 1, 4, 8, 8), torch.zeros(1, 1, 1, 1, 8, 8),
                                         torch.eye(3).expand(1, 1, 1, 3, 3), torch.eye(4)[:3].expand(1, 1, 1, 3, 4))
    with pytest.raises(_lib.MvtError):
        ops.camera_prep(torch.eye(3)[None], torch.eye(4)[:3][None])


def test_block_asserts_match_reference():
    """Same assertion conditions as mvtracker.py:782-788 (checked before any device work)."""
    x = torch.zeros(1, 10, 3)
    f = torch.zeros(1, 10, 6)
    with pytest.raises(AssertionError):
        dropin.PointcloudCorrBlock(11, 1, x, f, True, False)          # k > N
    with pytest.raises(AssertionError):
        dropin.PointcloudCorrBlock(4, 4, x, f, True, False)           # C % groups != 0
    with pytest.raises(AssertionError):
        dropin.PointcloudCorrBlock(4, 1, x, f, True, False, filter_invalid=True, valid=None)
    with pytest.raises(AssertionError):
        dropin.init_pointcloud_from_rgbd(torch.zeros(1, 1, 1, 4, 8, 8), torch.zeros(1, 1, 1, 1, 8, 9),
                                         torch.eye(3).expand(1, 1, 1, 3, 3), torch.eye(4)[:3].expand(1, 1, 1, 3, 4))


@pytest.mark.skipif(not ref_import.available(), reason="reference checkout not present (GPU box)")
def test_install_rebinds_reference_seam():
    ref = ref_import.load()
    saved = (ref.mvt.init_pointcloud_from_rgbd, ref.mvt.knn, ref.mvt.PointcloudCorrBlock)
    undo = dropin.install(ref.mvt)
    try:
        assert ref.mvt.init_pointcloud_from_rgbd is dropin.init_pointcloud_from_rgbd
        assert ref.mvt.knn is dropin.knn
        assert ref.mvt.PointcloudCorrBlock is dropin.PointcloudCorrBlock
        import inspect
        for ours, theirs in zip((dropin.init_pointcloud_from_rgbd, dropin.knn), saved[:2]):
            assert list(inspect.signature(ours).parameters) == list(inspect.signature(theirs).parameters)
        assert list(inspect.signature(dropin.PointcloudCorrBlock.__init__).parameters) == \
            list(inspect.signature(saved[2].__init__).parameters)
        assert list(inspect.signature(dropin.PointcloudCorrBlock.corr_sample).parameters) == \
            list(inspect.signature(saved[2].corr_sample).parameters)
    finally:
        undo()
    assert (ref.mvt.init_pointcloud_from_rgbd, ref.mvt.knn, ref.mvt.PointcloudCorrBlock) == saved


def test_adjacent_entry_points_check_arguments_without_a_gpu():
    """Rows f1-f3: the new entry points reject bad arguments before any device work (code -1 + message)."""
    L = _lib.lib()
    d = ctypes.c_void_p(256)
    assert L.mvt_track_input(d, d, 0, d, d, None, None, 12, 4, 128, 63, 256, d, 581 * 12, 581, 1, None) == -1       # odd E
    assert b"even" in L.mvt_last_error()
    assert L.mvt_track_input(d, d, 0, d, d, None, None, 65, 4, 128, 64, 256, d, 581 * 65, 581, 1, None) == -1       # S > 64
    assert L.mvt_track_input(d, None, 0, d, d, None, None, 12, 4, 128, 64, 256, d, 581 * 12, 581, 1, None) == -1
    assert b"null" in L.mvt_last_error()
    assert L.mvt_knn_corr_backward(d, d, 0, d, d, 64, 0, 1, 10, 5, 6, 1, 4, 4, d, 0, d, None) == -1                  # C % 4
    assert b"multiple of 4" in L.mvt_last_error()
    assert L.mvt_knn_corr_backward(d, d, 0, d, d, 64, 0, 1, 10, 5, 8, 1, 64, 4, d, 0, d, None) == -1                 # k > 32
    assert L.mvt_cloud_build_backward(d, 2, 2, 8, 16, 16, 4, d, 0, None) == -1                                      # level > 3
    assert L.mvt_cloud_build_backward(d, 2, 2, 8, 4, 16, 3, d, 0, None) == -1                                       # empty level
    assert b"empty" in L.mvt_last_error()
    assert L.mvt_lift_pixels(None, 4, 4, d, d, 3, 0, 0.0, d, None) == -1
    assert L.mvt_lift_pixels(d, 4, 4, d, d, 0, 0, 0.0, d, None) == 0                                                 # nothing to do


def test_torch_library_ops_have_fake_kernels_and_no_cpu_kernel():
    """torch.ops.mvt.* (mvtracker_b200/torch_ops.py): shapes / dtypes from the fake kernels (what torch.compile / export
    trace with); CPU tensors are refused by the dispatcher -- no CPU implementation exists."""
    import pytest
    import torch
    from torch._subclasses.fake_tensor import FakeTensorMode
    import mvtracker_b200.torch_ops  # noqa: F401  (registers the ops)
    with FakeTensorMode():
        fm, dp = torch.empty(3, 12, 128, 40, 56, device="cuda"), torch.empty(3, 12, 40, 56, device="cuda")
        K, E = torch.empty(3, 12, 3, 3, device="cuda"), torch.empty(3, 12, 3, 4, device="cuda")
        outs = torch.ops.mvt.cloud_build(fm, dp, K, E, 4.0, [0, 1, 2, 3], False)
        assert [tuple(o.shape) for o in outs[:3]] == [(12, 6720, 3), (12, 6720, 128), (12, 6720, 4)] and outs[-1].shape == (36, 48)
        q, t = torch.empty(12, 100, 3, device="cuda"), torch.empty(12, 100, 128, device="cuda")
        d, i = torch.ops.mvt.knn(outs[2], outs[-1], q, 16, 3, 40, 56, 4.0)
        assert d.shape == (12, 100, 16) and i.dtype == torch.int64
        c = torch.ops.mvt.knn_corr(outs[2::3][:4], outs[1::3][:4], outs[-1], t, q, 16, 3, 40, 56, 4.0)
        assert c.shape == (12, 100, 256) and c.dtype == torch.float32
        x = torch.ops.mvt.track_input(q, t, torch.empty(12, 100, device="cuda"), torch.empty(12, 100, device="cuda"),
                                      torch.empty(100, 581, device="cuda"), torch.empty(12, 581, device="cuda"), c, 64)
        assert x.shape == (100, 12, 581)
    with pytest.raises(NotImplementedError):
        torch.ops.mvt.knn(torch.zeros(1, 10, 4), torch.zeros(1, 48), torch.zeros(1, 2, 3), 1, 1, 2, 5, 4.0)
