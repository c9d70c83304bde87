#!/usr/bin/env python
"""Experiment builds: scripts/build_variants.py tag1:-DFOO=1,-DBAR=2 tag2: ...  ->  p-ntfields_b200/pntf/variants/libpntf_b200.<tag>.so
scripts/run_variants.sh then swaps each into place on the GPU box and runs a short bench."""
import importlib.util, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b", os.path.join(ROOT, "p-ntfields_b200", "build.py"))
b = importlib.util.module_from_spec(spec); spec.loader.exec_module(b)
vd = os.path.join(ROOT, "p-ntfields_b200", "pntf", "variants")
os.makedirs(vd, exist_ok=True)
for arg in sys.argv[1:]:
    tag, _, defs = arg.partition(":")
    defs = [d for d in defs.split(",") if d]
    print(tag, defs, b.build(force=True, defs=defs, out=os.path.join(vd, "libpntf_b200.%s.so" % tag), objdir="build_" + tag))
