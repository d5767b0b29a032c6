#!/usr/bin/env python
"""Copy the reference's own test fixtures for the render hot path into tests/golden/.

Run in the build container (where /root/reference exists); the copies are committed so
that the tests can run on the GPU box, where the reference tree is absent.

Sources (all under /root/reference/visr_bear/test/files):
  tensorfile_{le,be}_{1,32}.tenf      TENF reader KATs      (test_tensorfile.cpp:39-61)
  unity_brirs_decorrelators.tf        path-delay fixture    (test_delays.py:114-136)
  input.npy, *_out*.npy               end-to-end goldens    (test_changes.py:294-361; rendered
                                      by the reference with default.tf, which is NOT in the tree)
"""
import hashlib
import json
import shutil
from pathlib import Path

SRC = Path("/root/reference/visr_bear/test/files")
DST = Path(__file__).resolve().parent.parent / "tests" / "golden"


def main():
    DST.mkdir(parents=True, exist_ok=True)
    manifest = {}
    for p in sorted(SRC.iterdir()):
        if p.suffix in (".tenf", ".tf", ".npy"):
            shutil.copyfile(p, DST / p.name)
            manifest[p.name] = hashlib.sha256(p.read_bytes()).hexdigest()
    (DST / "MANIFEST.json").write_text(json.dumps(manifest, indent=1, sort_keys=True) + "\n")
    print("copied", len(manifest), "fixtures to", DST)


if __name__ == "__main__":
    main()
