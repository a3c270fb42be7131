"""Recipe for ``oracle/_ref``: the reference's pure-Python package packed into ONE zip archive, so that the GPU box (where
``/root/reference`` does not exist) can time and test against the reference's OWN solver classes.

    python oracle/make_ref.py            # build container only; does nothing when /root/reference is absent

``oracle/_ref/`` is git-ignored (reference sources never enter this repository's history) but not gpurun-ignored, so the
archive travels with the snapshot like the built ``.so``.  The reference has no build step (pure Python; a ``pip install
--target`` of it fails on pbr's git-metadata requirement and would add nothing), so "building" it is packing
``pycsou/**/*.py`` unmodified; ``oracle/ref_shim.py`` imports it straight from the archive (zipimport).
"""
from __future__ import annotations

import os
import sys
import zipfile

SRC = '/root/reference/pycsou'
ARCHIVE = os.path.join(os.path.dirname(os.path.abspath(__file__)), '_ref', 'pycsou_reference.zip')


def make_ref(verbose: bool = False) -> bool:
    """Create / refresh the archive from the live tree; returns whether an archive exists afterwards."""
    if not os.path.isdir(SRC):
        return os.path.exists(ARCHIVE)
    files = []
    for root, dirs, names in os.walk(SRC):
        dirs[:] = sorted(d for d in dirs if d != '__pycache__')
        files += [os.path.join(root, f) for f in sorted(names) if f.endswith('.py')]
    newest = max(os.path.getmtime(f) for f in files)
    if os.path.exists(ARCHIVE) and os.path.getmtime(ARCHIVE) >= newest:
        return True
    os.makedirs(os.path.dirname(ARCHIVE), exist_ok=True)
    tmp = ARCHIVE + '.tmp'
    with zipfile.ZipFile(tmp, 'w', zipfile.ZIP_DEFLATED) as z:
        for f in files:
            z.write(f, os.path.join('pycsou', os.path.relpath(f, SRC)))
    os.replace(tmp, ARCHIVE)
    if verbose:
        print(f'oracle/_ref: {len(files)} reference files -> {ARCHIVE}')
    return True


if __name__ == '__main__':
    sys.exit(0 if make_ref(verbose=True) else 1)
