"""Import the REAL reference (read-only, /root/reference) in the build container.

TEST INFRASTRUCTURE ONLY; used by tests/golden/make_golden.py to produce the
committed golden vectors, and by CPU tests that are skipped when
/root/reference is absent (it never exists on the GPU box).

Stubs: `easydict` (model_utils.py:10) and `png` (datasets/utils.py:14) are not
installed; `mvtracker.datasets/__init__.py` pulls matplotlib, so a bare package
object with only __path__ stands in for it.  Nothing of the reference is copied.
"""
import os
import sys
import types

REF_ROOT = os.environ.get("MVT_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "mvtracker"))


_cached = None


def load():
    """-> namespace with .mvt (the mvtracker.py module), .init_pointcloud_from_rgbd, .knn, .PointcloudCorrBlock"""
    global _cached
    if _cached is not None:
        return _cached
    if not available():
        raise RuntimeError(f"reference not found under {REF_ROOT}")
    sys.dont_write_bytecode = True
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    if "easydict" not in sys.modules:
        m = types.ModuleType("easydict")

        class EasyDict(dict):
            __getattr__ = dict.get
            __setattr__ = dict.__setitem__

        m.EasyDict = EasyDict
        sys.modules["easydict"] = m
    if "png" not in sys.modules:
        sys.modules["png"] = types.ModuleType("png")
    import mvtracker  # noqa: the reference package

    if "mvtracker.datasets" not in sys.modules:
        ds = types.ModuleType("mvtracker.datasets")
        ds.__path__ = [os.path.join(os.path.dirname(mvtracker.__file__), "datasets")]
        sys.modules["mvtracker.datasets"] = ds
    import logging

    logging.disable(logging.WARNING)
    try:
        import mvtracker.models.core.mvtracker.mvtracker as mvt
        from mvtracker.models.core.model_utils import init_pointcloud_from_rgbd
    finally:
        logging.disable(logging.NOTSET)
    ns = types.SimpleNamespace(mvt=mvt, init_pointcloud_from_rgbd=init_pointcloud_from_rgbd, knn=mvt.knn,
                               knn_torch=mvt._knn_torch, PointcloudCorrBlock=mvt.PointcloudCorrBlock,
                               MVTracker=mvt.MVTracker)
    _cached = ns
    return ns
