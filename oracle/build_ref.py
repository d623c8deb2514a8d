"""TEST INFRASTRUCTURE ONLY.  Recipe that packs the reference's OWN Python sources for the hot path into one
binary archive, `oracle/_ref/exoground_reference.pyz`, straight from where they lie under /root/reference.

Why: the GPU box has no /root/reference.  The archive is git-ignored (it never enters the history) but is not
gpurun-ignored, so it travels with the snapshot exactly like the built `libexg.so`.  On the box it lets
  * the `-m gpu` parity tests compare the CUDA path with the LIVE, UNMODIFIED reference modules on the same device
    (tests/test_live_reference_gpu.py) at the sizes BASELINE.json names,
  * `bench.py --impl reference` time the reference's real CPU code (`cpu_baseline.kind == "reference"`),
  * `bench.py` record the PyTorch-eager bar of the reference modules on the same B200 (`gpu_eager_baseline`).
Nothing under `exoground_b200/` imports it (tests/test_abi_cpu.py checks that).  No reference source is copied into
the tracked tree: the archive is a build product, imported through `zipimport` (oracle/ref_import.py).

    python -m oracle.build_ref        # no-op (exit 0) when /root/reference is absent and the archive exists
"""
import hashlib
import os
import sys
import zipfile

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_ref")
ARCHIVE = os.path.join(OUT_DIR, "exoground_reference.pyz")
REF = os.environ.get("EXG_REFERENCE_ROOT", "/root/reference")

# directories of the reference whose *.py files the path, its callers (trainers, alignment eval) and
# their import chains need.  Data files, notebooks, images and checkpoints are not packed.
DIRS = ("model", "train", "utils", "eval", "data", "end2end")


def _files():
    for d in DIRS:
        root = os.path.join(REF, d)
        for dp, _, fns in sorted(os.walk(root)):
            for fn in sorted(fns):
                if fn.endswith(".py"):
                    full = os.path.join(dp, fn)
                    yield full, os.path.relpath(full, REF)


def build(verbose=True):
    if not os.path.isdir(os.path.join(REF, "model")):
        if verbose:
            state = "present" if os.path.exists(ARCHIVE) else "ABSENT"
            print(f"oracle/_ref: no reference tree at {REF}; archive {state}")
        return ARCHIVE if os.path.exists(ARCHIVE) else None
    os.makedirs(OUT_DIR, exist_ok=True)
    files = list(_files())
    h = hashlib.sha256()
    for full, rel in files:
        h.update(rel.encode())
        with open(full, "rb") as f:
            h.update(f.read())
    digest = h.hexdigest()
    stamp = os.path.join(OUT_DIR, "exoground_reference.sha256")
    if os.path.exists(ARCHIVE) and os.path.exists(stamp) and open(stamp).read().strip() == digest:
        if verbose:
            print(f"oracle/_ref: up to date ({len(files)} files) -> {ARCHIVE}")
        return ARCHIVE
    tmp = ARCHIVE + ".tmp"
    with zipfile.ZipFile(tmp, "w", zipfile.ZIP_DEFLATED) as z:
        # explicit directory entries: zipimport treats a directory without __init__.py as a namespace package only
        # when the archive lists it (the reference's utils/, data/, eval/ have no __init__.py)
        dirs = set()
        for _, rel in files:
            d = os.path.dirname(rel)
            while d:
                dirs.add(d)
                d = os.path.dirname(d)
        for d in sorted(dirs):
            z.writestr(zipfile.ZipInfo(d + "/", date_time=(2020, 1, 1, 0, 0, 0)), b"")
        for full, rel in files:
            zi = zipfile.ZipInfo(rel, date_time=(2020, 1, 1, 0, 0, 0))        # reproducible archive
            zi.compress_type = zipfile.ZIP_DEFLATED
            with open(full, "rb") as f:
                z.writestr(zi, f.read())
    os.replace(tmp, ARCHIVE)
    with open(stamp, "w") as f:
        f.write(digest + "\n")
    if verbose:
        print(f"oracle/_ref: packed {len(files)} reference files -> {ARCHIVE}")
    return ARCHIVE


if __name__ == "__main__":
    build()
    sys.exit(0)
