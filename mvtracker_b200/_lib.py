"""ctypes binding of libmvt.so (the C ABI in include/mvt.h).  No CPU fallback: if the library is missing the import
of any op fails loudly (build it with `python -c "import __graft_entry__ as g; g.build()"` or `make -C mvtracker_b200/csrc`)."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("MVT_LIB") or os.path.join(_HERE, "libmvt.so")   # MVT_LIB: A/B builds in experiments

MVT_F32, MVT_BF16 = 0, 1
MAX_LEVELS, MAX_K, MAX_JOBS = 4, 32, 8
CAM_STRIDE, MAX_VIEWS = 48, 32
MVT_OK, MVT_E_ARG, MVT_E_CUDA, MVT_E_WORKSPACE = 0, -1, -2, -3
NOALLOC_BYTES = 4 << 20                          # MVT_KNN_CORR_NOALLOC_BYTES: static scratch of mvt_knn_corr_noalloc
NOALLOC_QUERIES = (NOALLOC_BYTES - 256) // 16    # MVT_KNN_CORR_NOALLOC_QUERIES

SYMBOLS = ("mvt_version", "mvt_last_error", "mvt_camera_prep", "mvt_cloud_build", "mvt_cloud_build_frames", "mvt_grid_workspace_bytes",
           "mvt_grid_build", "mvt_knn", "mvt_knn_corr", "mvt_knn_corr_workspace_bytes", "mvt_knn_corr_noalloc", "mvt_knn_corr_shared", "mvt_knn_gather", "mvt_grid_info", "mvt_launch_count", "mvt_debug_set_stats",
           "mvt_debug_set_group_width", "mvt_knn_keys", "mvt_topk_merge", "mvt_corr_from_idx",
           "mvt_cloud_build_shared", "mvt_knn_idx_shared", "mvt_corr_from_idx_shared", "mvt_tile_depth_floats",
           "mvt_tile_depth_build", "mvt_track_input", "mvt_knn_corr_backward",
           "mvt_cloud_build_backward", "mvt_lift_pixels", "mvt_shard_iter", "mvt_shard_local_bytes")


class LevelOut(C.Structure):
    _fields_ = [("xyz", C.c_void_p), ("xyz4", C.c_void_p), ("fvec", C.c_void_p), ("valid", C.c_void_p),
                ("spacing", C.c_void_p), ("neg_depth", C.c_void_p)]


class GridJob(C.Structure):
    _fields_ = [("xyz", C.c_void_p), ("mask", C.c_void_p), ("spacing", C.c_void_p), ("workspace", C.c_void_p),
                ("B", C.c_int), ("N", C.c_int), ("k_hint", C.c_int), ("index_mode", C.c_int), ("cell_scale", C.c_float)]


class Search(C.Structure):
    _fields_ = [("grid", C.c_void_p), ("xyz", C.c_void_p), ("xyz4", C.c_void_p), ("cam", C.c_void_p),
                ("cam_view_stride", C.c_int), ("N", C.c_int), ("V", C.c_int), ("H", C.c_int), ("W", C.c_int),
                ("stride", C.c_float), ("skip_invalid", C.c_int), ("tile_z", C.c_void_p), ("neg_depth", C.c_void_p)]


class CorrLevel(C.Structure):
    _fields_ = [("search", Search), ("fvec", C.c_void_p), ("k", C.c_int), ("out_offset", C.c_longlong)]


class OwnedLevel(C.Structure):
    _fields_ = [("xyz4", C.c_void_p), ("fvec", C.c_void_p), ("lo", C.c_longlong), ("hi", C.c_longlong),
                ("out_offset", C.c_longlong)]


class Peers(C.Structure):
    _fields_ = [("mc", C.c_void_p), ("ptr", C.c_void_p * 8), ("n", C.c_int)]


class ShardLevel(C.Structure):
    _fields_ = [("search", Search), ("xyz4_local", C.c_void_p), ("fvec_local", C.c_void_p), ("n_local", C.c_longlong),
                ("hw", C.c_int)]


class ShardCtx(C.Structure):
    _fields_ = [("inbox", Peers), ("ctrl", Peers), ("result", Peers), ("local", C.c_void_p), ("inbox_cap", C.c_longlong),
                ("rank", C.c_int), ("n_ranks", C.c_int), ("view_first", C.c_int * 9)]


class MvtError(RuntimeError):
    pass


_lib = None


def build(force=False):
    """Compile libmvt.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-s"] + (["-B"] if force else [])
    subprocess.check_call(cmd)
    return SO_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise MvtError(f"{SO_PATH} is missing: the CUDA library has not been built (there is no CPU fallback). "
                       f"Run `make -C {os.path.join(_HERE, 'csrc')}`.")
    L = C.CDLL(SO_PATH)
    vp, ci, cf, ll = C.c_void_p, C.c_int, C.c_float, C.c_longlong
    L.mvt_version.restype = ci
    L.mvt_last_error.restype = C.c_char_p
    L.mvt_launch_count.restype = C.c_ulonglong
    L.mvt_camera_prep.argtypes = [vp, vp, ci, vp, vp]
    L.mvt_cloud_build.argtypes = [vp, ci, vp, vp, ci, ci, ci, ci, ci, cf, ci, C.POINTER(LevelOut), ci, vp]
    L.mvt_cloud_build_frames.argtypes = [vp, ll, ci, vp, ll, vp, ci, ci, ci, ci, ci, ci, cf, ci, C.POINTER(LevelOut), ci, C.POINTER(Peers), ci, ci, vp]
    L.mvt_grid_workspace_bytes.argtypes = [ci, ci]
    L.mvt_grid_workspace_bytes.restype = C.c_size_t
    L.mvt_grid_build.argtypes = [C.POINTER(GridJob), ci, vp]
    L.mvt_knn.argtypes = [C.POINTER(Search), ci, vp, ci, ci, vp, ci, vp, vp]
    L.mvt_knn_corr.argtypes = [C.POINTER(CorrLevel), ci, ci, ci, ci, ci, ci, vp, vp, ci, ci, vp, ll, ll, vp, vp, vp, ci, vp]
    L.mvt_knn_corr_workspace_bytes.argtypes = [ci, ci, ci]
    L.mvt_knn_corr_shared.argtypes = [C.POINTER(CorrLevel), ci, ci, ci, ci, ci, vp, vp, ll, C.POINTER(Peers), ll, ll, vp, ci, vp]
    L.mvt_knn_corr_noalloc.argtypes = [C.POINTER(CorrLevel), ci, ci, ci, ci, ci, ci, vp, vp, ci, ci, vp, ll, ll, vp, vp, ci, vp]
    L.mvt_knn_corr_workspace_bytes.restype = C.c_size_t
    L.mvt_knn_gather.argtypes = [C.POINTER(Search), vp, ci, ci, ci, vp, ci, vp, vp, vp]
    L.mvt_grid_info.argtypes = [vp, ci, ci, vp, vp]
    L.mvt_debug_set_stats.argtypes = [vp]
    L.mvt_debug_set_group_width.argtypes = [ci]
    L.mvt_knn_keys.argtypes = [C.POINTER(Search), C.POINTER(ll), ci, ci, vp, ci, ci, ci, vp, ci, vp, vp, vp, vp]
    L.mvt_cloud_build_shared.argtypes = [vp, ci, vp, vp, ci, ci, ci, ci, ci, cf, ci, C.POINTER(LevelOut), ci, C.POINTER(Peers), ci, ci, vp]
    L.mvt_knn_idx_shared.argtypes = [C.POINTER(Search), ci, ci, vp, ci, ci, ci, ci, C.POINTER(Peers), vp]
    L.mvt_corr_from_idx_shared.argtypes = [C.POINTER(OwnedLevel), ci, ci, ci, ci, ci, ci, ci, ci, vp, vp, vp, C.POINTER(Peers), ll,
                                           C.POINTER(vp), C.POINTER(ll), vp]
    L.mvt_shard_iter.argtypes = [C.POINTER(ShardLevel), ci, ci, ci, ci, ci, ci, vp, vp, C.POINTER(ShardCtx), ll, C.c_ulonglong, vp]
    L.mvt_shard_local_bytes.argtypes = [ci, ci, ci, ci]
    L.mvt_shard_local_bytes.restype = C.c_size_t
    L.mvt_topk_merge.argtypes = [vp, ci, ll, ci, vp, vp, vp]
    L.mvt_corr_from_idx.argtypes = [C.POINTER(OwnedLevel), ci, ci, ci, ci, ci, ci, ci, vp, vp, vp, vp, ll, vp]
    L.mvt_tile_depth_floats.argtypes = [ci, ci, ci, ci]
    L.mvt_tile_depth_floats.restype = ll
    L.mvt_tile_depth_build.argtypes = [C.POINTER(Search), ci, ci, C.POINTER(vp), vp]
    L.mvt_track_input.argtypes = [vp, vp, ci, vp, vp, vp, vp, ci, ci, ci, ci, ci, vp, ll, ll, ci, vp]
    L.mvt_knn_corr_backward.argtypes = [vp, vp, ci, vp, vp, ll, ll, ci, ci, ci, ci, ci, ci, ci, vp, ci, vp, vp]
    L.mvt_lift_pixels.argtypes = [vp, ci, ci, vp, vp, ci, ci, cf, vp, vp]
    L.mvt_cloud_build_backward.argtypes = [vp, ci, ci, ci, ci, ci, ci, vp, ci, vp]
    for name in SYMBOLS:
        f = getattr(L, name)
        if name not in ("mvt_last_error", "mvt_grid_workspace_bytes", "mvt_launch_count", "mvt_tile_depth_floats",
                        "mvt_knn_corr_workspace_bytes", "mvt_shard_local_bytes"):
            f.restype = ci
    _lib = L
    return L


def check(rc, what):
    if rc != 0:
        msg = lib().mvt_last_error().decode("utf-8", "replace")
        raise MvtError(f"{what} failed (code {rc}): {msg}")
