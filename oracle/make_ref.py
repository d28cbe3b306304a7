#!/usr/bin/env python
"""Recipe for oracle/_ref/: the reference's OWN Python modules of the hot path, taken unmodified from
/root/reference/src so that they travel to the GPU box with the working tree (oracle/_ref/ is git-ignored, like the
built .so files, and is not gpurun-ignored).  bench.py times these modules as the CPU baseline (`kind: "reference"`) and
as the same-box eager-GPU comparator; tests/golden/make_golden.py imports the same files from /root/reference directly.

    python oracle/make_ref.py            # copies when /root/reference exists; otherwise reports what is there

Nothing in the product path (newsrecommendation_b200/) imports oracle/_ref; no reference source is committed."""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = '/root/reference/src'
OUT = os.path.join(HERE, '_ref')
FILES = ['model/__init__.py', 'model/NRMS.py', 'model/NAML.py', 'model/model_utils.py', 'metrics.py']


def make(verbose=True):
    if not os.path.isdir(REF_SRC):
        if verbose:
            print(f'oracle/make_ref.py: {REF_SRC} absent (GPU box); using the prebuilt oracle/_ref: '
                  f'{"present" if available() else "ABSENT"}')
        return available()
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(REF_SRC, rel), os.path.join(OUT, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        manifest[rel] = hashlib.sha256(open(dst, 'rb').read()).hexdigest()
    json.dump({'source': REF_SRC, 'sha256': manifest}, open(os.path.join(OUT, 'MANIFEST.json'), 'w'), indent=1)
    if verbose:
        print(f'oracle/make_ref.py: copied {len(FILES)} unmodified reference files into {OUT}')
    return True


def available():
    return all(os.path.exists(os.path.join(OUT, rel)) for rel in FILES)


def load():
    """(NRMS module, NAML module, metrics module) of the unmodified reference, or None when oracle/_ref is absent.
    The reference's package is called `model` (src/main.py:63: importlib.import_module(f'model.{args.model}'))."""
    if not available():
        return None
    if OUT not in sys.path:
        sys.path.insert(0, OUT)
    import importlib
    return (importlib.import_module('model.NRMS'), importlib.import_module('model.NAML'),
            importlib.import_module('metrics'))


if __name__ == '__main__':
    sys.exit(0 if make() else 1)
