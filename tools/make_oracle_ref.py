#!/usr/bin/env python
"""Copies the reference's own implementation of the hot path, UNMODIFIED, into oracle/_ref/ (git-ignored, so the
sources never enter this repo's history; gpurun-shipped, so the GPU box can time them):

    /root/reference/TorchMD_Fix/forces.py   -> oracle/_ref/forces.py     (Forces.compute, forces.py:82-344)
    /root/reference/TorchMD_Fix/systems.py  -> oracle/_ref/systems.py    (System, systems.py:5-85)

plus oracle/_ref/MANIFEST.json with their sha256.  `bench.py --impl reference` and the cpu_baseline leg import them
by path (oracle/ref_arm.py) when present and report kind "reference"; without them they time the torch port and say
"port".  Runs only where /root/reference exists (this container); __graft_entry__.build() calls it.

    python tools/make_oracle_ref.py [--ref /root/reference]
"""
import argparse
import hashlib
import json
import os
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FILES = ("forces.py", "systems.py")


def make(ref="/root/reference"):
    src = os.path.join(ref, "TorchMD_Fix")
    if not all(os.path.exists(os.path.join(src, f)) for f in FILES):
        return False
    dst = os.path.join(ROOT, "oracle", "_ref")
    os.makedirs(dst, exist_ok=True)
    manifest = {"source": src, "files": {}}
    for f in FILES:
        shutil.copyfile(os.path.join(src, f), os.path.join(dst, f))
        with open(os.path.join(dst, f), "rb") as fh:
            manifest["files"][f] = hashlib.sha256(fh.read()).hexdigest()
    with open(os.path.join(dst, "MANIFEST.json"), "w") as fh:
        json.dump(manifest, fh, indent=1)
    return True


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    ok = make(ap.parse_args().ref)
    print("oracle/_ref written" if ok else "reference not found: nothing written")
