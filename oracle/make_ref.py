#!/usr/bin/env python
"""Recipe: stage the UNMODIFIED reference modules under oracle/_ref/ (git-ignored, shipped to the
GPU box with the snapshot like the built .so files).

    python oracle/make_ref.py [/root/reference]

The reference is 12 plain Python files with no build step (no setup.py / pyproject), so "building"
it is copying the files where they lie, byte for byte; a manifest with their SHA-256 is written
next to them.  Nothing under oracle/_ref is committed and nothing in the product imports it.
Users: tests/test_reference_scripts_gpu.py (runs the reference's own experiment scripts on the
B200 modules with only PYTHONPATH changed) and bench.py's CPU arm (times the reference's own
train.train_epoch on the host cores, cpu_baseline.kind = "reference").
__graft_entry__.build() runs this when /root/reference exists (the build container only).
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")


def stage(src="/root/reference"):
    if not os.path.isdir(src):
        return None
    os.makedirs(DEST, exist_ok=True)
    manifest = {}
    for name in sorted(os.listdir(src)):
        path = os.path.join(src, name)
        if os.path.isfile(path) and name.endswith((".py", ".txt")):
            shutil.copyfile(path, os.path.join(DEST, name))
            with open(path, "rb") as f:
                manifest[name] = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(DEST, "MANIFEST.json"), "w") as f:
        json.dump({"source": src, "sha256": manifest}, f, indent=1, sort_keys=True)
    return manifest


if __name__ == "__main__":
    m = stage(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
    print("no reference tree: nothing staged" if m is None else f"staged {len(m)} files under {DEST}")
