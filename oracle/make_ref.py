"""Stage the UNMODIFIED reference files of the hot path's host model into oracle/_ref/ (git-ignored, not gpurun-ignored).

TEST / BASELINE INFRASTRUCTURE ONLY.  /root/reference does not exist on the GPU box; what this script stages travels there
with the repo snapshot exactly like the built .so files do, so that
  * tests/test_gpu_model.py can run the reference's own MVTracker.forward on the B200 with the stock seam and with
    mvtracker_b200.dropin.install(), and compare traj_e / vis_e,
  * bench.py --impl reference can time the reference's OWN init_pointcloud_from_rgbd / PointcloudCorrBlock.corr_sample
    (-> _knn_torch, mvtracker/models/core/mvtracker/mvtracker.py:75-79) on the box's host cores,
  * bench.py can report end-to-end tracked frames/s of MVTracker.forward with and without the drop-in.
Nothing below is committed: the files are byte-for-byte copies made at build time (the reference has no setup.py /
pyproject, so `pip install --target` cannot produce them) and oracle/_ref/ is listed in .gitignore.  The product package
(mvtracker_b200/) never imports from here.

The staged subset is the import closure of mvtracker.models.core.mvtracker.mvtracker:
    mvtracker/models/core/mvtracker/mvtracker.py   MVTracker, PointcloudCorrBlock, knn
    mvtracker/models/core/model_utils.py           init_pointcloud_from_rgbd, smart_cat, bilinear_sample2d
    mvtracker/models/core/embeddings.py            sin / cos embeddings
    mvtracker/models/core/cotracker2/blocks.py     EfficientUpdateFormer
    mvtracker/models/core/spatracker/blocks.py     BasicEncoder
    mvtracker/datasets/utils.py                    transform_scene (imported at module level by mvtracker.py:12)
    mvtracker/utils/basic.py                       homogeneous helpers
plus the packages' __init__.py files (mvtracker/datasets/__init__.py is NOT staged: it imports the dataset classes and
with them matplotlib; oracle/ref_import.py installs a bare package object for it, as it does for /root/reference).

    python oracle/make_ref.py            # stage (no-op when /root/reference is absent and _ref already exists)
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("MVT_REFERENCE_ROOT", "/root/reference")
DEST = os.path.join(HERE, "_ref")

FILES = [
    "mvtracker/__init__.py",
    "mvtracker/models/__init__.py",
    "mvtracker/models/core/__init__.py",
    "mvtracker/models/core/mvtracker/__init__.py",
    "mvtracker/models/core/mvtracker/mvtracker.py",
    "mvtracker/models/core/model_utils.py",
    "mvtracker/models/core/embeddings.py",
    "mvtracker/models/core/cotracker2/__init__.py",
    "mvtracker/models/core/cotracker2/blocks.py",
    "mvtracker/models/core/spatracker/__init__.py",
    "mvtracker/models/core/spatracker/blocks.py",
    "mvtracker/datasets/utils.py",
    "mvtracker/utils/__init__.py",
    "mvtracker/utils/basic.py",
]


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def stage(verbose=True) -> bool:
    """-> True when oracle/_ref is usable afterwards."""
    if not os.path.isdir(os.path.join(REF_ROOT, "mvtracker")):
        ok = os.path.isfile(os.path.join(DEST, "MANIFEST.json"))
        if verbose:
            print(f"make_ref: {REF_ROOT} absent; " + ("keeping the staged copy" if ok else "nothing staged"))
        return ok
    manifest = {}
    for rel in FILES:
        src = os.path.join(REF_ROOT, rel)
        if not os.path.isfile(src):
            raise FileNotFoundError(src)
        dst = os.path.join(DEST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        manifest[rel] = _sha(dst)
    with open(os.path.join(DEST, "MANIFEST.json"), "w") as f:
        json.dump({"source": REF_ROOT, "files": manifest}, f, indent=1, sort_keys=True)
    if verbose:
        print(f"make_ref: staged {len(manifest)} unmodified reference files into {DEST}")
    return True


def verify() -> bool:
    """The staged files are byte-identical to what MANIFEST.json recorded (i.e. nobody edited the 'reference')."""
    mf = os.path.join(DEST, "MANIFEST.json")
    if not os.path.isfile(mf):
        return False
    files = json.load(open(mf))["files"]
    return all(os.path.isfile(os.path.join(DEST, rel)) and _sha(os.path.join(DEST, rel)) == h for rel, h in files.items())


if __name__ == "__main__":
    sys.exit(0 if stage() else 1)
