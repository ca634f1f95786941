#!/bin/bash
# TEST INFRASTRUCTURE ONLY (oracle). Compiles the reference's own native smoother /
# setup kernels (SWIG wrapper amg_core_wrap.cxx + header templates relaxation.h, graph.h,
# smoothed_aggregation.h, ruge_stuben.h, ...) FROM WHERE THEY LIE under /root/reference
# into oracle/_ref/_amg_core.so.  Nothing is copied; only the built .so lands in oracle/_ref/
# (git-ignored, travels to the GPU box).  Flags follow pyamg/amg_core/setup.py:12-14.
set -e
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
mkdir -p "$OUT"
if [ ! -d "$REF/pyamg/amg_core" ]; then
  echo "build_ref: $REF absent; keeping prebuilt $OUT/_amg_core.so" >&2
  exit 0
fi
if [ "$OUT/_amg_core.so" -nt "$REF/pyamg/amg_core/amg_core_wrap.cxx" ] && [ -z "$FORCE" ]; then
  exit 0
fi
PYINC=$(python3 -c "import sysconfig;print(sysconfig.get_paths()['include'])")
NPINC=$(python3 -c "import numpy;print(numpy.get_include())")
g++ -O3 -std=c++14 -fPIC -shared -w -D__STDC_FORMAT_MACROS=1 \
    -I"$PYINC" -I"$NPINC" -I"$REF/pyamg/amg_core" \
    "$REF/pyamg/amg_core/amg_core_wrap.cxx" -o "$OUT/_amg_core.so"
echo "built $OUT/_amg_core.so"
