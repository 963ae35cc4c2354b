"""Tuning builds of the CUDA core: the same library with different -D macros on some translation units.
    python tools/build_variants.py name=DEF1,DEF2[@src1.cu+src2.cu] ...
Outputs simple-path-tracer_b200/lib/variants/libspt_<name>.so (git-ignored; travels to the GPU box)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

for spec in sys.argv[1:]:
    name, rest = spec.split("=", 1)
    only = None
    if "@" in rest:
        rest, srcs = rest.split("@")
        only = srcs.split("+")
    defs = [d for d in rest.split(",") if d]
    out = os.path.join(g.PKG, "lib", "variants", f"libspt_{name}.so")
    g.build_lib(out, defs=defs, only=only)
    print("built", out, defs, only)
