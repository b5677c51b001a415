#!/usr/bin/env python
"""Stage the UNMODIFIED reference (disk02/ComfyUI-SeedVR2) for the GPU box:  python scripts/stage_reference.py

/root/reference exists only in the build container and is read-only; the GPU box receives a snapshot of this repository.
baseline/_ref/ is git-ignored (reference sources never enter the history) but NOT gpurun-ignored, so what is copied here
travels with the snapshot.  The reference is a ComfyUI custom node, not an installable package (its top-level __init__
needs comfy_api; `pip install --target baseline/_ref /root/reference` has nothing to build), so "installing" it is copying its
pure-Python `src/` tree and the two model YAMLs:

    baseline/_ref/src/...            the reference's own modules, byte for byte (sha256 manifest in baseline/_ref/MANIFEST.json)
    baseline/_ref/configs_{3b,7b}/   main.yaml

Users: tests/test_gpu_dropin.py and tests/test_gpu_fullwidth.py (the REAL NaDiT / NaMMSRTransformerBlock on the B200),
bench.py --impl reference and bench.py's GPU-vs-GPU baseline leg — all through oracle/refstage.py, never the product.
Third-party imports the reference needs and the image lacks (rotary_embedding_torch, diffusers, omegaconf) come from
oracle/shim, as in tests/golden/make_golden.py."""
import hashlib
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get("SVR2_REFERENCE", "/root/reference")
DST = os.path.join(ROOT, "baseline", "_ref")
WHAT = ["src", "configs_3b", "configs_7b"]


def main() -> int:
    if not os.path.isdir(os.path.join(SRC, "src", "models")):
        print(f"stage_reference: {SRC} not present — nothing staged (the GPU box uses what the build container staged)")
        return 0
    manifest = {}
    for name in WHAT:
        d = os.path.join(DST, name)
        if os.path.isdir(d):
            shutil.rmtree(d)
        shutil.copytree(os.path.join(SRC, name), d, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
        for base, _, files in os.walk(d):
            for f in sorted(files):
                p = os.path.join(base, f)
                with open(p, "rb") as fh:
                    manifest[os.path.relpath(p, DST)] = hashlib.sha256(fh.read()).hexdigest()
    with open(os.path.join(DST, "MANIFEST.json"), "w") as f:
        json.dump({"source": SRC, "files": manifest}, f, indent=0, sort_keys=True)
    print(f"stage_reference: {len(manifest)} files -> {DST}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
