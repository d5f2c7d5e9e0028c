"""Stage the UNMODIFIED reference's hot-path modules into oracle/_ref/ (git-ignored, NOT gpurun-ignored).

Test / measurement infrastructure, not product code.  The reference is pure Python, so its "build" is a file
copy: this recipe imports the reference from /root/reference through oracle/ref_shims.py (in-process shims
only, no edits), looks at which of its source files the sampling path actually loaded, and copies exactly
those files byte for byte -- with their package layout -- into ONE archive, oracle/_ref/reference_path.zip, which
Python imports directly (zipimport; a sys.path entry may be a zip file).  Nothing under oracle/_ref/ enters
the git history (.gitignore) but it travels to the GPU box with the snapshot, like a built .so, so that
`bench.py`'s `cpu_baseline` / `--impl reference` legs can time the reference's own CPU code there
(`cpu_baseline.kind = "reference"`).  Run in the build container only:

    python oracle/stage_ref.py

`__graft_entry__.build()` calls `stage()` when /root/reference is present.  A manifest with the sha256 of
every staged file is written next to the archive; `verify()` re-checks it (tests/test_oracle_golden.py).
"""
import hashlib
import json
import os
import shutil
import sys
import zipfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
DEST = os.path.join(HERE, "_ref")
ARCHIVE = "reference_path.zip"
SRC = "/root/reference"


def _sha(path):
    with open(path, "rb") as fh:
        return hashlib.sha256(fh.read()).hexdigest()


def stage(src=SRC, dest=DEST):
    """Archive the reference modules the sampling path imports under `dest`.  Returns the manifest dict."""
    if not os.path.isdir(os.path.join(src, "diffusion")):
        raise RuntimeError("reference tree not present at %s" % src)
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    os.environ["GD_REFERENCE_ROOT"] = src
    from oracle import ref_shims
    if ref_shims.REF_ROOT != src:
        raise RuntimeError("ref_shims already bound to %s" % ref_shims.REF_ROOT)
    ref_shims.load_reference()
    prefix = os.path.join(os.path.realpath(src), "")
    files = set()
    for mod in list(sys.modules.values()):
        f = getattr(mod, "__file__", None)
        if f and os.path.realpath(f).startswith(prefix) and f.endswith(".py"):
            files.add(os.path.realpath(f))
    if os.path.isdir(dest):
        shutil.rmtree(dest)
    os.makedirs(dest)
    manifest = {}
    with zipfile.ZipFile(os.path.join(dest, ARCHIVE), "w", zipfile.ZIP_DEFLATED) as z:
        dirs = set()
        for f in sorted(files):
            rel = os.path.relpath(f, prefix)
            d = os.path.dirname(rel)
            while d and d not in dirs:                      # explicit directory entries: the reference's packages have no
                dirs.add(d)                                 # __init__.py (namespace packages), zipimport finds those by entry
                z.writestr(d + "/", "")
                d = os.path.dirname(d)
            z.write(f, rel)
            manifest[rel] = _sha(f)
    with open(os.path.join(dest, "MANIFEST.json"), "w") as fh:
        json.dump({"source": src, "archive": ARCHIVE, "files": manifest}, fh, indent=1, sort_keys=True)
    return manifest


def verify(dest=DEST, src=None):
    """True when every archived file still has its manifest digest (and equals the source file when `src` exists)."""
    path = os.path.join(dest, "MANIFEST.json")
    if not os.path.exists(path) or not os.path.exists(os.path.join(dest, ARCHIVE)):
        return False
    man = json.load(open(path))["files"]
    with zipfile.ZipFile(os.path.join(dest, ARCHIVE)) as z:
        for rel, dig in man.items():
            if hashlib.sha256(z.read(rel)).hexdigest() != dig:
                return False
            if src and os.path.exists(os.path.join(src, rel)) and _sha(os.path.join(src, rel)) != dig:
                return False
    return bool(man)


if __name__ == "__main__":
    m = stage()
    print("staged %d reference files into %s" % (len(m), DEST))
