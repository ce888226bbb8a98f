"""Stage the reference's own experiment drivers for the GPU box -- TEST INFRASTRUCTURE, not product.

The drop-in claim is "code/setups/*.py run unchanged against this package".  Proving it means executing those
scripts, byte for byte, on a B200 -- but /root/reference does not exist there, and reference sources must never
enter this repository's history.  So, like a compiled reference under oracle/_ref/, the scripts are COPIED AS
BUILD OUTPUTS into oracle/_ref/setups/ (git-ignored, shipped with the working tree by gpurun) whenever the
reference checkout is present at build time.  Nothing is edited; tests/test_gpu_unmodified_setups.py checks the
SHA-256 of every staged file against tests/golden/setups_sha256.json before running it.

    python oracle/stage_ref.py            # stage + (re)write the digest manifest
"""
import glob
import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SETUPS = "/root/reference/code/setups"
STAGED = os.path.join(HERE, "_ref", "setups")
MANIFEST = os.path.join(os.path.dirname(HERE), "tests", "golden", "setups_sha256.json")


def sha256(path):
    with open(path, "rb") as fh:
        return hashlib.sha256(fh.read()).hexdigest()


def stage(write_manifest=False):
    """Copy code/setups/*.py into oracle/_ref/setups/ when the reference is present; returns the staged paths."""
    if not os.path.isdir(REF_SETUPS):
        return sorted(glob.glob(os.path.join(STAGED, "*.py")))
    os.makedirs(STAGED, exist_ok=True)
    digests = {}
    for src in sorted(glob.glob(os.path.join(REF_SETUPS, "*.py"))):
        dst = os.path.join(STAGED, os.path.basename(src))
        shutil.copyfile(src, dst)
        digests[os.path.basename(src)] = sha256(dst)
    if write_manifest:
        with open(MANIFEST, "w") as fh:
            json.dump(digests, fh, indent=1, sort_keys=True)
            fh.write("\n")
    return sorted(glob.glob(os.path.join(STAGED, "*.py")))


if __name__ == "__main__":
    for p in stage(write_manifest=True):
        print(p)
