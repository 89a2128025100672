"""Recipe: stage the reference's own PyTorch implementation of the hot path under oracle/_ref/.

TEST / BASELINE INFRASTRUCTURE ONLY (never imported by the product package).

The reference is a Python script collection (no setup.py / pyproject, nothing to compile), so "building" it
means placing the unmodified package `nerf-pytorch/nerf/` (run_one_iter_of_nerf and everything it calls:
train_utils.py, nerf_helpers.py, volume_rendering_utils.py, models.py, cfgnode.py + the loaders its __init__
imports) where the GPU box can import it: /root/reference does not exist there, oracle/_ref/ travels with the
gpurun snapshot.  oracle/_ref/ is git-ignored: reference SOURCES never enter the history; this script does.

    python oracle/build_ref.py            # copies when /root/reference is present, verifies otherwise

Used by: bench.py --impl reference, bench.py's cpu_baseline leg (kind "reference"), tests that pin the oracle.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("NERF_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")
# (source relative to the reference root, destination relative to oracle/_ref)
TREES = [("nerf-pytorch/nerf", "nerf-pytorch/nerf")]
FILES = [("parser_instant_ngp.py", "parser_instant_ngp.py")]


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def build(verbose=False):
    """Copy (when the reference tree is here) and return the manifest {relative path: sha256}; {} if neither the
    reference nor a previous copy exists."""
    manifest_path = os.path.join(DST, "MANIFEST.json")
    if os.path.isdir(os.path.join(REF_SRC, "nerf-pytorch", "nerf")):
        os.makedirs(DST, exist_ok=True)
        manifest = {}
        for src, dst in TREES:
            s, d = os.path.join(REF_SRC, src), os.path.join(DST, dst)
            if os.path.isdir(d):
                shutil.rmtree(d)
            shutil.copytree(s, d, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
            for root, _, names in os.walk(d):
                for n in sorted(names):
                    p = os.path.join(root, n)
                    rel = os.path.relpath(p, DST)
                    manifest[rel] = _sha(p)
                    assert manifest[rel] == _sha(os.path.join(REF_SRC, src, os.path.relpath(p, d))), rel  # unmodified
        for src, dst in FILES:
            s, d = os.path.join(REF_SRC, src), os.path.join(DST, dst)
            if os.path.exists(s):
                shutil.copyfile(s, d)
                manifest[dst] = _sha(d)
        with open(manifest_path, "w") as f:
            json.dump({"source": REF_SRC, "files": manifest}, f, indent=1, sort_keys=True)
        if verbose:
            print("oracle/_ref: %d files staged from %s" % (len(manifest), REF_SRC))
        return manifest
    if os.path.exists(manifest_path):
        manifest = json.load(open(manifest_path))["files"]
        for rel, sha in manifest.items():
            assert _sha(os.path.join(DST, rel)) == sha, "oracle/_ref/%s was modified" % rel
        if verbose:
            print("oracle/_ref: %d files verified against MANIFEST.json" % len(manifest))
        return manifest
    return {}


if __name__ == "__main__":
    m = build(verbose=True)
    sys.exit(0 if m else 1)
