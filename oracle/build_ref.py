"""Recipe for oracle/_ref: the UNMODIFIED reference files, copied where they lie under
/root/reference into oracle/_ref/ so that they travel to the GPU box (oracle/_ref/ is
git-ignored: no reference source enters the history; it is not gpurun-ignored).

THIS IS TEST INFRASTRUCTURE.  Nothing under pytorch-dppo_b200/ imports it; bench.py's
``--impl reference`` / ``cpu_baseline`` legs run it through oracle/ref_runner.py.

    python oracle/build_ref.py

The reference is Python: there is nothing to compile and no build system to run.  The files
are byte-for-byte copies (sha256 recorded in oracle/_ref/MANIFEST.json)."""
import hashlib
import json
import os
import shutil

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
FILES = ("model.py", "train.py", "chief.py", "utils.py", "ppo.py", "main.py", "test.py")


def build():
    if not os.path.isdir(REF):
        return os.path.isdir(OUT)                     # on the GPU box: use what travelled
    os.makedirs(OUT, exist_ok=True)
    manifest = {}
    for f in FILES:
        src = os.path.join(REF, f)
        shutil.copyfile(src, os.path.join(OUT, f))
        manifest[f] = hashlib.sha256(open(src, "rb").read()).hexdigest()
    with open(os.path.join(OUT, "MANIFEST.json"), "w") as fh:
        json.dump({"source": REF, "sha256": manifest}, fh, indent=1)
    return True


if __name__ == "__main__":
    print("oracle/_ref:", "ready" if build() else "unavailable (no /root/reference here)")
