#!/usr/bin/env python
"""Recipe for oracle/_ref/: the UNMODIFIED reference modules of the hot path, taken from where they lie.

The reference is pure Python (no build): when /root/reference is present (the development container), the four
files `models/{hash_encoding,encoders,configs,transformers}.py` are copied byte for byte into oracle/_ref/models/ so
that the GPU box -- which has no /root/reference -- can run the reference's own classes as the CPU baseline
(`bench.py --impl reference --workload c1r`) and as the PyTorch-CUDA eager comparator.  oracle/_ref/ is git-ignored
(never part of the repo's history) and is test / baseline infrastructure: nothing under deepearth_b200/ may load it.

    python oracle/make_ref.py          (also run by __graft_entry__.build())
"""
import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("DE_REFERENCE_ROOT", "/root/reference")
FILES = ["hash_encoding.py", "encoders.py", "configs.py", "transformers.py", "cross_modal_fusion.py"]


def make_ref(verbose: bool = False) -> bool:
    """Copies the files if the reference tree is there; returns whether oracle/_ref is usable afterwards."""
    dst = os.path.join(HERE, "_ref", "models")
    src = os.path.join(REF_ROOT, "models")
    if os.path.isdir(src):
        os.makedirs(dst, exist_ok=True)
        manifest = []
        for f in FILES:
            shutil.copyfile(os.path.join(src, f), os.path.join(dst, f))
            with open(os.path.join(dst, f), "rb") as fh:
                manifest.append(f"{hashlib.sha256(fh.read()).hexdigest()}  models/{f}")
        with open(os.path.join(HERE, "_ref", "MANIFEST.sha256"), "w") as fh:
            fh.write("\n".join(manifest) + "\n")
        if verbose:
            print("\n".join(manifest))
    return all(os.path.exists(os.path.join(dst, f)) for f in FILES)


if __name__ == "__main__":
    ok = make_ref(verbose=True)
    print("oracle/_ref", "ready" if ok else "unavailable (no reference tree)")
    sys.exit(0)
