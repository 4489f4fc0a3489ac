#!/usr/bin/env bash
# Links baseline/harness.cu and the reference's pyngp module against the libraries baseline/build_ref.sh made.
#   baseline/_ref/ref_harness                      tcnn / NerfNetwork kernel-grain driver (baseline/harness.cu)
#   baseline/_ref/pyngp.cpython-312-*.so           the reference's own Python binding (src/python_api.cu), headless
# The binding is compiled against the venv's pybind11 (3.x): the vendored pybind11 2.10 predates Python 3.12.
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=${OUT:-$HERE/_ref}
ARCH=${ARCH:-100}
TCNN=$REF/dependencies/tiny-cuda-nn
[ -d "$REF/src" ] || { echo "build_harness: $REF absent, keeping prebuilt files" >&2; exit 0; }
if [ "$ARCH" = "90" ]; then GENCODE="-gencode arch=compute_90,code=compute_90"; else GENCODE="-gencode arch=compute_${ARCH},code=sm_${ARCH}"; fi
COMMON="-std=c++14 -O3 -DNDEBUG --extended-lambda --expt-relaxed-constexpr -Xcompiler=-fPIC,-fno-strict-aliasing,-Wno-float-conversion -Xcudafe=--diag_suppress=unrecognized_gcc_pragma -w -DTCNN_MIN_GPU_ARCH=$ARCH $GENCODE"
TCNN_INC="-I$TCNN/include -I$TCNN/dependencies -I$TCNN/dependencies/cutlass/include -I$TCNN/dependencies/cutlass/tools/util/include -I$TCNN/dependencies/fmt/include"
NGP_INC="-I$REF/include -I$REF/dependencies -I$REF/dependencies/filesystem -I$REF/dependencies/nanovdb -I$REF/dependencies/NaturalSort -I$REF/dependencies/tinylogger -I$REF/dependencies/zstr/src"
LIBS="-L/usr/local/cuda/lib64/stubs -lcuda -lz -ldl -lpthread"

if [ "$HERE/harness.cu" -nt "$OUT/ref_harness" ] || [ "$OUT/obj/libtcnn_ref.a" -nt "$OUT/ref_harness" ]; then
  nvcc $COMMON --use_fast_math $TCNN_INC $NGP_INC -c "$HERE/harness.cu" -o "$OUT/obj/harness.o"
  nvcc $GENCODE -o "$OUT/ref_harness" "$OUT/obj/harness.o" "$OUT/obj/libtcnn_ref.a" $LIBS
  echo "build_harness: $OUT/ref_harness" >&2
fi

PYSO="$OUT/pyngp$(python3 -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')"
if [ ! -f "$PYSO" ] || [ "$OUT/obj/libngp_ref.a" -nt "$PYSO" ]; then
  PYB=$(python3 -c "import pybind11;print(pybind11.get_include())")
  PYI=$(python3 -c "import sysconfig;print(sysconfig.get_paths()['include'])")
  [ -f "$OUT/obj/py_python_api.o" ] || nvcc $COMMON --use_fast_math -rdc=true -DNGP_VERSION='"1.0dev"' -DNGP_PYTHON -Xcompiler=-fvisibility=hidden \
    -I"$PYB" -I"$PYI" $TCNN_INC $NGP_INC -c "$REF/src/python_api.cu" -o "$OUT/obj/py_python_api.o"
  # device-link the binding together with the ngp objects (separable compilation, CMakeLists.txt:331)
  NGP_OBJS=$(ls "$OUT"/obj/ngp_*.o | grep -v ngp_dlink.o)
  nvcc $GENCODE -shared -Xcompiler=-fPIC -o "$PYSO" "$OUT/obj/py_python_api.o" $NGP_OBJS "$OUT/obj/libtcnn_ref.a" $LIBS
  echo "build_harness: $PYSO" >&2
fi
