"""TEST INFRASTRUCTURE ONLY - builds the UNMODIFIED reference hot path into ``oracle/_ref/``.

The reference's hot path is Python, so "building it from the sources where they lie" means byte-compiling the
files under ``/root/reference`` (``compile`` + ``marshal``, same interpreter as the GPU box: same image) into
``oracle/_ref/<package>/<module>.refbin`` (a neutral extension: the snapshot sent to the GPU box drops ``*.pyc``);
``oracle/ref_harness.py`` installs an importer for them.  No reference source is copied into the repository; ``oracle/_ref/``
is git-ignored but NOT gpurun-ignored, so the compiled modules travel to the GPU box like our own ``.so``.
``__graft_entry__.build()`` runs this when ``/root/reference`` is present (i.e. in the build container).

Used by: ``oracle/ref_harness.py`` (falls back to ``oracle/_ref`` when ``/root/reference`` is absent), and through
it by ``bench.py --impl reference`` / the ``cpu_baseline`` leg (``kind: "reference"``) and by the GPU parity tests
that compare against the reference itself instead of the port.
"""
from __future__ import annotations

import json
import marshal
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REFERENCE_ROOT = "/root/reference"
# the files on the path of SURVEY.md 8(a), plus what they import at module level
FILES = ["reconstruct/__init__.py", "reconstruct/optimizer.py", "reconstruct/loss.py", "reconstruct/loss_utils.py",
         "reconstruct/utils.py", "reconstruct/tools.py",
         "deep_sdf/__init__.py", "deep_sdf/deep_sdf_decoder.py", "deep_sdf/workspace.py"]


def build(verbose: bool = True) -> str | None:
    if not os.path.isfile(os.path.join(REFERENCE_ROOT, FILES[1])):
        if verbose:
            print("oracle/build_ref: /root/reference is absent, keeping", OUT if os.path.isdir(OUT) else "(nothing)")
        return OUT if os.path.isdir(OUT) else None
    manifest = {"python": sys.version.split()[0], "files": {}}
    for rel in FILES:
        src = os.path.join(REFERENCE_ROOT, rel)
        dst = os.path.join(OUT, rel[:-3] + ".refbin")            # X.py -> X.refbin: the marshalled module code object
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        with open(src, "rb") as f:
            code = compile(f.read(), rel, "exec", dont_inherit=True)
        with open(dst, "wb") as f:
            marshal.dump(code, f)
        manifest["files"][rel] = os.path.getsize(src)
    with open(os.path.join(OUT, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1)
    if verbose:
        print("oracle/build_ref: compiled", len(FILES), "reference modules into", OUT)
    return OUT


if __name__ == "__main__":
    build()
