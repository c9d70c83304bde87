"""Recipe for oracle/_ref: the reference's own implementation of the hot path, made available to `bench.py --impl reference`.

TEST / MEASUREMENT INFRASTRUCTURE ONLY (like everything under oracle/).  The reference is pure Python: there is nothing
to compile.  In the build container (where /root/reference exists) this copies, unmodified and byte for byte, the files
of the reference that define its `models` package for this path

    models/__init__.py, models/model_res_sigmoid.py, models/model_res_sigmoid_multi.py,
    models/data_mlp.py, models/data_multi.py            (the two loaders the model files import at module top)

into oracle/_ref/models/.  oracle/_ref/ is git-ignored (reference sources never enter the history) but not
gpurun-ignored, so it travels to the GPU box, where /root/reference does not exist.  `load()` imports that copy through
the same 2-module shim the golden generator uses (empty matplotlib / matplotlib.pylab, pickle5 -> pickle).
"""
import hashlib
import os
import pickle
import shutil
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
REF_DST = os.path.join(HERE, "_ref")
FILES = ("models/__init__.py", "models/model_res_sigmoid.py", "models/model_res_sigmoid_multi.py", "models/data_mlp.py",
         "models/data_multi.py")


def make(verbose=False):
    """Copies the files when the reference checkout is present; returns True when oracle/_ref is usable."""
    if os.path.isdir(REF_SRC):
        for f in FILES:
            dst = os.path.join(REF_DST, f)
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            shutil.copyfile(os.path.join(REF_SRC, f), dst)
        with open(os.path.join(REF_DST, "MANIFEST.txt"), "w") as m:
            for f in FILES:
                h = hashlib.sha256(open(os.path.join(REF_DST, f), "rb").read()).hexdigest()
                m.write("%s  %s\n" % (h, f))
        if verbose:
            print("oracle/_ref: copied %d files from %s" % (len(FILES), REF_SRC))
    return available()


def available():
    return all(os.path.exists(os.path.join(REF_DST, f)) for f in FILES)


def load():
    """(model_res_sigmoid, model_res_sigmoid_multi) of the UNMODIFIED reference, imported from oracle/_ref."""
    if not available():
        raise RuntimeError("oracle/_ref is missing: run `python oracle/make_ref.py` in the build container")
    for name in ("matplotlib", "matplotlib.pylab"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules.setdefault("pickle5", pickle)
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "models" or k.startswith("models.")}
    sys.path.insert(0, REF_DST)
    try:
        from models import model_res_sigmoid as S, model_res_sigmoid_multi as M
    finally:
        sys.path.remove(REF_DST)
        for k in list(sys.modules):
            if k == "models" or k.startswith("models."):
                del sys.modules[k]
        sys.modules.update(saved)
    return S, M


if __name__ == "__main__":
    print("available:", make(verbose=True))
