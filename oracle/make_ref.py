"""TEST INFRASTRUCTURE ONLY -- recipe that stages the UNMODIFIED reference package for the
timed CPU arm (`bench.py --impl reference`, `cpu_baseline.kind == "reference"`).

    python -m oracle.make_ref            # /root/reference/src/garage -> oracle/_ref/garage

The reference is pure Python (no build step): "building" it is a byte-for-byte copy of its
package directory from where it lies under /root/reference into the git-ignored `oracle/_ref/`
(listed in .gitignore, NOT in .gpurunignore, so it travels to the GPU box like libb2g.so).
Nothing is copied into tracked files; nothing under garage_b200/ reads `oracle/_ref/`.
`oracle/ref_shims.py` resolves `import garage` to /root/reference/src when that exists (build
container) and to `oracle/_ref` otherwise (GPU box); the third-party imports the reference
makes at package-import time (dowel, gym, akro, ...) are stubbed there.
A manifest (file count, total bytes, sha256 over the sorted file list) is written beside the
copy so that a run can state which tree it timed.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = '/root/reference/src/garage'
DST_ROOT = os.path.join(HERE, '_ref')
DST = os.path.join(DST_ROOT, 'garage')


def manifest(root):
    h = hashlib.sha256()
    n = total = 0
    for d, _, files in sorted(os.walk(root)):
        for f in sorted(files):
            if not f.endswith('.py'):
                continue
            p = os.path.join(d, f)
            with open(p, 'rb') as fh:
                b = fh.read()
            h.update(os.path.relpath(p, root).encode())
            h.update(b)
            n += 1
            total += len(b)
    return dict(files=n, bytes=total, sha256=h.hexdigest())


def make(force=False):
    """Copy the package if the source exists; returns the manifest or None when neither the
    source nor an earlier copy is present."""
    if os.path.isdir(SRC):
        want = manifest(SRC)
        mf = os.path.join(DST_ROOT, 'MANIFEST.json')
        if not force and os.path.exists(mf):
            try:
                if json.load(open(mf)) == want:
                    return want
            except ValueError:
                pass
        if os.path.isdir(DST):
            shutil.rmtree(DST)
        os.makedirs(DST_ROOT, exist_ok=True)
        shutil.copytree(SRC, DST, ignore=shutil.ignore_patterns('__pycache__', '*.pyc'))
        version = '/root/reference/VERSION'
        if os.path.exists(version):
            shutil.copy(version, os.path.join(DST_ROOT, 'VERSION'))
        with open(mf, 'w') as fh:
            json.dump(want, fh)
        return want
    mf = os.path.join(DST_ROOT, 'MANIFEST.json')
    if os.path.exists(mf):
        return json.load(open(mf))
    return None


if __name__ == '__main__':
    m = make(force='--force' in sys.argv)
    print('oracle/_ref:', m if m else 'reference source not present and no earlier copy')
