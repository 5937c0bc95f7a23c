#!/usr/bin/env python
"""Recipe for ``oracle/_ref``: the reference's OWN rl_step, made importable on the GPU box.

TEST / BENCH INFRASTRUCTURE ONLY -- nothing under pipelinerl_swe_b200/ imports this.

The reference is pure Python; its hot path (pipelinerl/finetune/rl/__init__.py:130-377 ``rl_step``, :380-501
``populate_rl_data``, rl/utils.py, finetune/types.py) needs only torch / numpy / pandas / pydantic / datasets /
transformers, all present in this image.  /root/reference does not exist on the GPU box, so this script stages those
four files, byte for byte, from where they lie under /root/reference into ``oracle/_ref/`` -- a build OUTPUT directory
that is git-ignored (reference sources never enter the history) but travels to the box with the working tree, exactly
like the built ``.so``.  ``__graft_entry__.build()`` runs it whenever /root/reference is present.

``load()`` imports the staged package and returns the reference's ``pipelinerl.finetune.rl`` module; bench.py uses it
as the CPU arm (``cpu_baseline.kind == "reference"``) and tests use it to re-pin the oracle restatement.
"""
from __future__ import annotations

import hashlib
import importlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("GRPO_REFERENCE", "/root/reference")
FILES = (
    "pipelinerl/finetune/__init__.py",
    "pipelinerl/finetune/types.py",
    "pipelinerl/finetune/rl/__init__.py",
    "pipelinerl/finetune/rl/utils.py",
)


def build(ref: str = REF, out: str = OUT) -> bool:
    """Stage the files; returns False (and leaves ``out`` alone) when the reference tree is not on this machine."""
    if not all(os.path.isfile(os.path.join(ref, f)) for f in FILES):
        return False
    manifest = {"source": ref, "files": {}}
    for f in FILES:
        dst = os.path.join(out, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(ref, f), dst)
        manifest["files"][f] = hashlib.sha256(open(dst, "rb").read()).hexdigest()
    with open(os.path.join(out, "MANIFEST.json"), "w") as fh:
        json.dump(manifest, fh, indent=1)
    return True


def available(out: str = OUT) -> bool:
    return all(os.path.isfile(os.path.join(out, f)) for f in FILES)


def load(out: str = OUT):
    """Import the staged reference and return its ``pipelinerl.finetune.rl`` module (raises ImportError if absent)."""
    if not available(out):
        raise ImportError(f"{out} is not staged: run `python oracle/build_ref.py` where /root/reference exists")
    if out not in sys.path:
        sys.path.insert(0, out)
    return importlib.import_module("pipelinerl.finetune.rl")


if __name__ == "__main__":
    ok = build()
    print(f"oracle/_ref: {'staged ' + str(len(FILES)) + ' files from ' + REF if ok else 'reference tree not found at ' + REF}")
    if ok:
        m = load()
        print("import ok:", m.rl_step.__module__, "rl_step")
