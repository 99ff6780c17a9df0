#!/usr/bin/env python
"""Build a tuning variant of libocsb200.so beside the product library:
    python tools/build_variant.py NAME [-DFLAG ...] [--maxrregcount=N ...]
-> build/variants/NAME.so (travels to the GPU box; load it with OCSB_LIB)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ocsmesh_b200 import build
name, extra = sys.argv[1], sys.argv[2:]
d = os.path.join(ROOT, "build", "variants")
os.makedirs(d, exist_ok=True)
print(build.build(extra=tuple(extra), lib=os.path.join(d, name + ".so"),
                  objdir=os.path.join(ROOT, "build", "obj_" + name)))
