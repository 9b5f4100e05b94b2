"""Recipe for ``oracle/_ref``: a byte-for-byte copy of the reference's hot-path Python package, so that the UNMODIFIED
reference implementation can run next to the CUDA path on the GPU box (where ``/root/reference`` does not exist).

    python -m oracle.build_ref            # needs /root/reference (build container); idempotent

TEST INFRASTRUCTURE ONLY.  ``oracle/_ref/`` is git-ignored (it never enters the history and no product module imports it)
but it is not gpurun-ignored, so it travels with the working tree exactly like the built ``libsam2b.so``.  Only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s reference / baseline legs load it, always as the checker or as
the thing being compared against, never as the product.

What is copied (nothing is edited; ``MANIFEST.json`` records the sha256 of every file so the tests can prove it):
  sam2/sam2/**/*.py, sam2/sam2/**/*.yaml     the model, predictors and configs (reference package ``sam2``)
  sam2/ytools/__init__.py, sam2/ytools/executor/*.py     the ``ModelExectuor`` plug-in ABC the predictors import
The TensorRT / ONNX Runtime executors, the native ``csrc`` and everything outside the package are left behind (they need
libraries that are not in this image and are not on the path).
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC_ROOT = os.environ.get("SAM2_REFERENCE_SRC", "/root/reference")
DST_ROOT = os.path.join(HERE, "_ref")
_KEEP_EXT = (".py", ".yaml")
_TREES = ("sam2/sam2", "sam2/ytools/executor")
_FILES = ("sam2/ytools/__init__.py",)


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as fh:
        h.update(fh.read())
    return h.hexdigest()


def _wanted():
    for tree in _TREES:
        top = os.path.join(SRC_ROOT, tree)
        for root, dirs, files in os.walk(top):
            dirs[:] = sorted(d for d in dirs if d not in ("__pycache__", "csrc"))
            for f in sorted(files):
                if f.endswith(_KEEP_EXT):
                    yield os.path.relpath(os.path.join(root, f), SRC_ROOT)
    for f in _FILES:
        yield f


def build(force=False):
    """Copy the package when the source tree is present; otherwise keep whatever is already there.  Returns the path of
    ``oracle/_ref`` if a usable copy exists afterwards, else None."""
    manifest_path = os.path.join(DST_ROOT, "MANIFEST.json")
    if not os.path.isdir(os.path.join(SRC_ROOT, "sam2", "sam2")):
        return DST_ROOT if os.path.exists(manifest_path) else None
    files = list(_wanted())
    want = {rel: _sha(os.path.join(SRC_ROOT, rel)) for rel in files}
    if not force and os.path.exists(manifest_path):
        have = json.load(open(manifest_path)).get("files", {})
        if have == want and all(os.path.exists(os.path.join(DST_ROOT, r)) for r in want):
            return DST_ROOT
    if os.path.isdir(DST_ROOT):
        shutil.rmtree(DST_ROOT)
    for rel in files:
        dst = os.path.join(DST_ROOT, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(SRC_ROOT, rel), dst)
    with open(manifest_path, "w") as fh:
        json.dump({"source": SRC_ROOT, "note": "unmodified copies; sha256 per file", "files": want}, fh, indent=1, sort_keys=True)
    return DST_ROOT


def verify():
    """True when every file under oracle/_ref still has the sha256 the manifest recorded at copy time."""
    manifest_path = os.path.join(DST_ROOT, "MANIFEST.json")
    if not os.path.exists(manifest_path):
        return False
    files = json.load(open(manifest_path))["files"]
    return all(os.path.exists(os.path.join(DST_ROOT, r)) and _sha(os.path.join(DST_ROOT, r)) == h for r, h in files.items())


if __name__ == "__main__":
    p = build(force="--force" in sys.argv)
    print(p, "verified" if p and verify() else "MISSING")
