#!/usr/bin/env python
"""Recipe for oracle/_ref/: the UNMODIFIED reference files the prototype-head path imports, placed where they can
travel to the GPU box (gpurun snapshots /root/repo only; /root/reference does not exist there).

TEST / BENCH INFRASTRUCTURE ONLY.  oracle/_ref/ is git-ignored (the reference's sources never enter this repository's
history) but not gpurun-ignored.  The reference is a script tree without packaging metadata (no setup.py /
pyproject.toml), so `pip install --target ... /root/reference` has nothing to install; this script copies, byte for
byte, the import closure of `model`, `segmentation.module`, `segmentation.push`, `find_nearest`, `prune` and
`train_and_test` (found by importing them here through oracle/ref_shims and listing sys.modules) and records a manifest with SHA-256 digests.

Used by:  bench.py --impl reference and bench.py's cpu_baseline leg (the reference's own
PatchClassificationModule._step on the box's host cores, oracle/ref_runner.py), __graft_entry__.build() (runs this
recipe when /root/reference is present).

    python oracle/build_ref.py            # (re)creates oracle/_ref/
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("ADLM_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")

# The import closure is found at build time: a child process imports the path's modules from the reference through the
# shims and lists every module file under the reference root.
_CLOSURE = r"""
import os, sys
ref, shims = sys.argv[1], sys.argv[2]
for k in ("DATA_PATH", "TRAIN_DIR", "TEST_DIR", "TRAIN_PUSH_DIR"):
    os.environ.setdefault(k, "/tmp/adlm_ref_data")
os.environ.setdefault("LOG_DIR", "/tmp/adlm_ref_log")
os.makedirs(os.environ["LOG_DIR"], exist_ok=True)
sys.path.insert(0, shims); sys.path.insert(0, ref)
import model, segmentation.module, segmentation.push, find_nearest, prune, train_and_test  # noqa
for m in list(sys.modules.values()):
    f = getattr(m, "__file__", None)
    if f and os.path.abspath(f).startswith(os.path.abspath(ref) + os.sep):
        print(os.path.relpath(os.path.abspath(f), os.path.abspath(ref)))
"""


def closure():
    import subprocess
    out = subprocess.check_output([sys.executable, "-c", _CLOSURE, REF, os.path.join(HERE, "ref_shims")], text=True,
                                  stderr=subprocess.DEVNULL)
    return sorted(set(l.strip() for l in out.splitlines() if l.strip().endswith(".py")))


def build(force=False):
    if not os.path.isdir(REF):
        return None
    manifest_path = os.path.join(OUT, "MANIFEST.json")
    if os.path.exists(manifest_path) and not force:
        return OUT
    if os.path.isdir(OUT):
        shutil.rmtree(OUT)
    manifest = {"source": REF, "files": {}}
    for rel in closure():
        src = os.path.join(REF, rel)
        if not os.path.exists(src):
            continue
        dst = os.path.join(OUT, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        manifest["files"][rel] = hashlib.sha256(open(src, "rb").read()).hexdigest()
    try:
        import subprocess
        manifest["commit"] = subprocess.check_output(["git", "-C", REF, "rev-parse", "HEAD"], text=True,
                                                     stderr=subprocess.DEVNULL).strip()
    except Exception:  # noqa: BLE001 - the reference may not be a git checkout
        manifest["commit"] = None
    with open(manifest_path, "w") as f:
        json.dump(manifest, f, indent=1)
    return OUT


if __name__ == "__main__":
    out = build(force=True)
    print(out if out else f"{REF} not present: nothing built", file=sys.stderr if not out else sys.stdout)
