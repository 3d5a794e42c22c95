"""Recipe for oracle/_ref: the UNMODIFIED reference package made importable on this image.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Run in the build container, where /root/reference
exists:  python oracle/make_ref.py   (also run by __graft_entry__.build()).

The reference is pure Python, so "building" it means: copy its `nums/` package, byte for byte, from
/root/reference into oracle/_ref/nums (git-ignored: nothing of the reference enters the history; the
directory travels to the GPU box with the snapshot like a built .so), and put next to it empty
stand-ins for the third-party packages this image lacks (ray, opt_einsum, boto3, botocore, zarr —
SURVEY.md section 8c).  `oracle/refharness.py` then restores the NumPy aliases NumPy 2 removed and
imports it.  A MANIFEST with the sha256 of every copied file is written so a test can check the
copy is untouched.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("NUMS_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "_ref")

STUBS = {
    "ray": "def is_initialized():\n    return False\n\n\ndef remote(*a, **k):\n    return lambda f: f\n",
    "opt_einsum": "import numpy as np\n\ncontract = np.einsum\n",
    "boto3": "",
    "botocore": "",
    "zarr": "",
}


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as fh:
        h.update(fh.read())
    return h.hexdigest()


def available():
    return os.path.isdir(os.path.join(REF, "nums"))


def build(force=False):
    """Creates oracle/_ref (if the reference is present).  Returns the path or None."""
    if not available():
        return OUT if os.path.isdir(os.path.join(OUT, "nums")) else None
    manifest_path = os.path.join(OUT, "MANIFEST.json")
    if not force and os.path.exists(manifest_path):
        return OUT
    if os.path.isdir(OUT):
        shutil.rmtree(OUT)
    os.makedirs(OUT)
    shutil.copytree(os.path.join(REF, "nums"), os.path.join(OUT, "nums"),
                    ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    for name, body in STUBS.items():
        os.makedirs(os.path.join(OUT, "_stubs", name))
        with open(os.path.join(OUT, "_stubs", name, "__init__.py"), "w") as fh:
            fh.write(body)
    os.makedirs(os.path.join(OUT, "data"), exist_ok=True)   # settings.py:22-23 creates <project_root>/data on import
    manifest = {}
    for base, _, files in os.walk(os.path.join(OUT, "nums")):
        for f in sorted(files):
            if f.endswith(".py"):
                p = os.path.join(base, f)
                manifest[os.path.relpath(p, OUT)] = _sha(p)
    with open(manifest_path, "w") as fh:
        json.dump({"source": REF, "files": manifest}, fh, indent=0, sort_keys=True)
    return OUT


if __name__ == "__main__":
    path = build(force="--force" in sys.argv)
    print(path if path else "reference not present at %s; nothing built" % REF)
