"""TEST / BASELINE INFRASTRUCTURE - not part of the product (nothing under deepfields-nerf-_b200/ imports this).

Stages the reference's OWN Python sources of the hot path under oracle/_ref/ so that `bench.py --impl reference` can time the
reference's unmodified `forward_pass` + backward + AdamW (scripts/train.py:129-186, :244-248, :294-323) on the GPU box's host
cores, where /root/reference does not exist.  oracle/_ref/ is git-ignored (the sources never enter the history) but not
gpurun-ignored, so it travels with the snapshot like the built .so files.

    python oracle/make_ref.py        # in the build container; __graft_entry__.build() runs it when /root/reference is present

Files staged, byte for byte: scripts/{train,model,sampling,dataset,instantngp}.py (train.py imports the other four at module
level).  Two third-party modules that are absent from this image are needed at import time only (omegaconf, tinycudann):
`load()` puts empty stand-ins into sys.modules, exactly as tests/golden/make_golden.py does for the golden vectors.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
SRC = "/root/reference/scripts"
FILES = ("train.py", "model.py", "sampling.py", "dataset.py", "instantngp.py")


def stage() -> bool:
    """Copies the files; returns False (and leaves oracle/_ref untouched) when the reference is not on this machine."""
    if not all(os.path.exists(os.path.join(SRC, f)) for f in FILES):
        return False
    os.makedirs(DST, exist_ok=True)
    manifest = {}
    for f in FILES:
        shutil.copyfile(os.path.join(SRC, f), os.path.join(DST, f))
        with open(os.path.join(DST, f), "rb") as fh:
            manifest[f] = hashlib.sha256(fh.read()).hexdigest()
    with open(os.path.join(DST, "MANIFEST.json"), "w") as fh:
        json.dump({"source": SRC, "sha256": manifest}, fh, indent=1)
    return True


def available() -> bool:
    return all(os.path.exists(os.path.join(DST, f)) for f in FILES)


def load():
    """Imports the staged reference train module (and returns it with model / sampling): `ref_train.forward_pass` etc."""
    if not available():
        raise RuntimeError("oracle/_ref is empty: run `python oracle/make_ref.py` where /root/reference exists")
    om = types.ModuleType("omegaconf")
    om.OmegaConf = type("OmegaConf", (), {})
    sys.modules.setdefault("omegaconf", om)
    sys.modules.setdefault("tinycudann", types.ModuleType("tinycudann"))
    sys.dont_write_bytecode = True
    if DST not in sys.path:
        sys.path.insert(0, DST)
    import model as ref_model          # noqa: E402  (the reference's module names)
    import sampling as ref_sampling    # noqa: E402
    import train as ref_train          # noqa: E402
    for m in (ref_model, ref_sampling, ref_train):
        assert os.path.dirname(os.path.abspath(m.__file__)) == DST, f"{m.__name__} resolved to {m.__file__}, not oracle/_ref"
    return ref_train, ref_model, ref_sampling


if __name__ == "__main__":
    print("staged" if stage() else "reference not found at " + SRC, "->", DST)
