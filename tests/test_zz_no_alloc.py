"""mvt_knn_corr_noalloc (include/mvt.h): the variant of mvt_knn_corr whose scratch -- the list of deferred large-box queries
-- is a static device array inside libmvt.so, for callers that cannot allocate.  It has no counterpart in the reference,
so the existing entry point mvt_knn_corr with a caller-owned workspace is
what it is checked against, bit for bit (the parity tests hold the fused kernels to bit-exact indices and identical floats).

Named test_zz_* so that pytest collects it last.
Both entry points are called through the C ABI (ctypes); torch only owns the input / output tensors.
"""
import ctypes as C

import pytest
import torch

from mvtracker_b200 import _lib, ops, synth

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _scene(dtype, seed, M, S=3):
    """Four views, four levels with tile tables; half of the tracks float in free space so that the deferred list is used."""
    sc = synth.make_scene(4, 80 * 4, 96 * 4, S=S, C=128, zero_frac=0.1, seed=seed, device=DEV, dtype=dtype)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 1, 2, 3], want_valid=True,
                             want_tiles=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], M, S, valid=clouds[0].valid[0], dtype=dtype)
    g = torch.Generator(device=DEV).manual_seed(seed + 1)
    coords[:, :M // 2] += 0.4 * torch.randn(S, M // 2, 3, device=DEV, generator=g)
    return clouds, coords.contiguous(), targets.contiguous()


def _levels(clouds, k):
    lv = (_lib.CorrLevel * _lib.MAX_LEVELS)()
    keep = []
    for i, c in enumerate(clouds):
        sr = ops._as_search(c)
        assert sr.c.tile_z, "the test needs tile tables: without them neither entry point touches its scratch"
        keep.append(sr)
        lv[i] = _lib.CorrLevel(sr.c, c.fvec.data_ptr(), k, i * k * 4)
    return lv, keep


def _ptr(t):
    return C.c_void_p(None if t is None else t.data_ptr())


def _call(entry, clouds, coords, targets, k, mode, workspace=None):
    """entry: 'mvt_knn_corr' (workspace: caller-owned uint8 tensor) or 'mvt_knn_corr_noalloc' -> (out, idx, return code)"""
    L = _lib.lib()
    B, M, Cc = targets.shape
    n = len(clouds)
    lv, keep = _levels(clouds, k)
    out = torch.full((B, M, n * k * 4), float("nan"), device=DEV, dtype=torch.float32)
    idx = torch.full((n, B, M, k), -7, device=DEV, dtype=torch.int32)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    head = (lv, n, B, M, Cc, 1, ops._feat_dtype(targets), _ptr(targets), _ptr(coords), 1, 0, _ptr(out), out.stride(1), out.stride(0),
            _ptr(None), _ptr(idx))
    if entry == "mvt_knn_corr":
        rc = L.mvt_knn_corr(*head, _ptr(workspace), mode, stream)
    else:
        rc = L.mvt_knn_corr_noalloc(*head, mode, stream)
    torch.cuda.synchronize()
    del keep
    return out, idx, rc


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_noalloc_matches_workspace_entry_point_twice(dtype):
    L = _lib.lib()
    k = 16
    # two different problems back to back: the second call finds the first call's list in the shared array
    for seed, M in ((95, 600), (31, 257)):
        clouds, coords, targets = _scene(dtype, seed, M)
        B = targets.shape[0]
        need = int(L.mvt_knn_corr_workspace_bytes(4, B, M))
        ws = torch.empty(need, device=DEV, dtype=torch.uint8)
        want, widx, rc = _call("mvt_knn_corr", clouds, coords, targets, k, 1, ws)
        assert rc == 0, L.mvt_last_error()
        deferred = ws[:8].view(torch.int32).cpu().tolist()
        assert deferred[0] > 50 and deferred[1] >= deferred[0], deferred        # the scratch really is in use on this input
        got, gidx, rc = _call("mvt_knn_corr_noalloc", clouds, coords, targets, k, 1)
        assert rc == 0, L.mvt_last_error()
        assert torch.equal(gidx, widx)
        assert torch.equal(got, want) and not torch.isnan(got).any()
        # the same call again, on the array the call before left dirty; and count-only mode (one kernel, same bits)
        got2, gidx2, rc = _call("mvt_knn_corr_noalloc", clouds, coords, targets, k, 1)
        assert rc == 0 and torch.equal(gidx2, widx) and torch.equal(got2, want)
        got3, gidx3, rc = _call("mvt_knn_corr_noalloc", clouds, coords, targets, k, 2)
        assert rc == 0 and torch.equal(gidx3, widx) and torch.equal(got3, want)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    L = _lib.lib()
    clouds, coords, targets = _scene(torch.float32, 95, 64, S=1)
    lv, keep = _levels(clouds, 16)
    out = torch.zeros(1, 64, 256, device=DEV)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    M_big = int(_lib.NOALLOC_QUERIES) // 4 + 1        # 4 levels, B = 1: one query too many
    # the shape is refused before anything is read or launched, so the 64-track tensors can stand in for the big ones
    n0 = L.mvt_launch_count()
    rc = L.mvt_knn_corr_noalloc(lv, 4, 1, M_big, 128, 1, _lib.MVT_F32, _ptr(targets), _ptr(coords), 1, 0, _ptr(out), 256, 0,
                                _ptr(None), _ptr(None), 1, stream)
    assert rc == _lib.MVT_E_WORKSPACE and b"exceeds the static scratch" in L.mvt_last_error()
    rc = L.mvt_knn_corr_noalloc(lv, 4, 1, 64, 128, 1, _lib.MVT_F32, _ptr(targets), _ptr(coords), 1, 0, _ptr(out), 256, 0,
                                _ptr(None), _ptr(None), 0, stream)
    assert rc == _lib.MVT_E_ARG and b"hard_query_mode" in L.mvt_last_error()
    assert L.mvt_launch_count() == n0
    # the largest accepted size is exactly what mvt_knn_corr_workspace_bytes says fits
    assert int(L.mvt_knn_corr_workspace_bytes(1, 1, int(_lib.NOALLOC_QUERIES))) == int(_lib.NOALLOC_BYTES)
    assert int(L.mvt_knn_corr_workspace_bytes(1, 1, int(_lib.NOALLOC_QUERIES) + 1)) > int(_lib.NOALLOC_BYTES)
    del keep
