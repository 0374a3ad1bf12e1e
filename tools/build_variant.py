"""Builds another copy of the core library with extra compile flags into build/variants/<name>/,
for A/B runs in one GPU session:  B200QUAD_LIB=build/variants/<name>/libb200quad.so python ...

  python tools/build_variant.py magic -DBQ_TRIG_MAGIC=1
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
name, flags = sys.argv[1], sys.argv[2:]
os.environ["BQ_NVCC_EXTRA"] = " ".join(flags)
import build  # noqa: E402  (reads BQ_NVCC_EXTRA at import)

build.LIBDIR = os.path.join(ROOT, "build", "variants", name)
build.OBJDIR = os.path.join(ROOT, "build", "variants", name + "_obj")
os.makedirs(build.LIBDIR, exist_ok=True)
os.makedirs(build.OBJDIR, exist_ok=True)
build.CXX_FLAGS = build.CXX_FLAGS + flags          # host objects see the same macros (host/device parity)
print(build.build_core(force=True))
