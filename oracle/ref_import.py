"""Import the REAL reference: from /root/reference in the build container, else from the unmodified files that
oracle/make_ref.py staged into oracle/_ref/ (git-ignored; travels to the GPU box like the built .so files).

TEST / BASELINE INFRASTRUCTURE ONLY; used by tests/golden/make_golden*.py to produce the committed golden vectors, by the
CPU tests that rebind the seam on the real module, by tests/test_gpu_model.py (reference forward vs drop-in forward on
the B200) and by bench.py's reference arm / end-to-end frames-per-second leg.  Never imported by mvtracker_b200/.

Stubs: `easydict` (model_utils.py:10) and `png` (datasets/utils.py:14) are not installed; `mvtracker.datasets/__init__.py`
pulls matplotlib, so a bare package object with only __path__ stands in for it.
"""
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
STAGED = os.path.join(HERE, "_ref")


def root():
    """Directory that holds the reference's `mvtracker` package, or None."""
    env = os.environ.get("MVT_REFERENCE_ROOT")
    for cand in ([env] if env else []) + ["/root/reference", STAGED]:
        if cand and os.path.isfile(os.path.join(cand, "mvtracker", "models", "core", "mvtracker", "mvtracker.py")):
            return cand
    return None


REF_ROOT = root() or "/root/reference"


def available():
    return root() is not None


def kind():
    """'reference' for /root/reference, 'staged' for oracle/_ref (the same files, copied unmodified by make_ref.py)."""
    r = root()
    return None if r is None else ("staged" if os.path.abspath(r) == os.path.abspath(STAGED) else "reference")


_cached = None


def load():
    """-> namespace with .mvt (the mvtracker.py module), .init_pointcloud_from_rgbd, .knn, .PointcloudCorrBlock, .MVTracker"""
    global _cached
    if _cached is not None:
        return _cached
    r = root()
    if r is None:
        raise RuntimeError("reference not found (neither /root/reference nor oracle/_ref; run `python oracle/make_ref.py` "
                           "where /root/reference exists)")
    sys.dont_write_bytecode = True
    if r not in sys.path:
        sys.path.insert(0, r)
    if "easydict" not in sys.modules:
        try:
            import easydict  # noqa: F401
        except Exception:
            m = types.ModuleType("easydict")

            class EasyDict(dict):
                __getattr__ = dict.get
                __setattr__ = dict.__setitem__

            m.EasyDict = EasyDict
            sys.modules["easydict"] = m
    if "png" not in sys.modules:
        try:
            import png  # noqa: F401
        except Exception:
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
                               MVTracker=mvt.MVTracker, root=r, kind=kind())
    _cached = ns
    return ns
