#!/usr/bin/env bash
# Builds integration/tcnn_caller.cpp - a program written against the reference's tcnn/include/tiny-cuda-nn/cpp_api.h only -
# twice:
#   integration/_build/tcnn_caller_b200   linked with integration/tcnn_cpp_api_b200.cpp + libngp_b200.so (this library)
#   integration/_build/tcnn_caller_ref    linked with the reference's own tcnn build (baseline/_ref/obj/libtcnn_ref.a)
# Needs the reference's headers, so it runs where /root/reference exists; the binaries travel to the GPU box.
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(dirname "$HERE")
OUT=$HERE/_build
TCNN=$REF/dependencies/tiny-cuda-nn
[ -d "$TCNN/include" ] || { echo "integration/build.sh: $REF absent, keeping prebuilt files" >&2; exit 0; }
mkdir -p "$OUT"
INC="-I$TCNN/include -I$TCNN/dependencies -I/usr/local/cuda/include -I$ROOT/include"
PKG=$ROOT/instant-ngp-rendering_b200
if [ ! -f "$OUT/tcnn_caller_b200" ] || [ "$HERE/tcnn_caller.cpp" -nt "$OUT/tcnn_caller_b200" ] || [ "$HERE/tcnn_cpp_api_b200.cpp" -nt "$OUT/tcnn_caller_b200" ] || [ "$ROOT/include/ngp_b200.h" -nt "$OUT/tcnn_caller_b200" ]; then
  g++ -std=c++14 -O2 -Wall $INC "$HERE/tcnn_caller.cpp" "$HERE/tcnn_cpp_api_b200.cpp" -L"$PKG" -lngp_b200 -Wl,-rpath,'$ORIGIN/../../instant-ngp-rendering_b200' \
    -L/usr/local/cuda/lib64 -lcudart -o "$OUT/tcnn_caller_b200"
  echo "integration/build.sh: $OUT/tcnn_caller_b200" >&2
fi
LIBREF=$ROOT/baseline/_ref/obj/libtcnn_ref.a
if [ -f "$LIBREF" ] && { [ ! -f "$OUT/tcnn_caller_ref" ] || [ "$HERE/tcnn_caller.cpp" -nt "$OUT/tcnn_caller_ref" ] || [ "$LIBREF" -nt "$OUT/tcnn_caller_ref" ]; }; then
  g++ -std=c++14 -O2 -Wall $INC -c "$HERE/tcnn_caller.cpp" -o "$OUT/tcnn_caller.o"
  nvcc -gencode arch=compute_100,code=sm_100 -o "$OUT/tcnn_caller_ref" "$OUT/tcnn_caller.o" "$LIBREF" -L/usr/local/cuda/lib64/stubs -lcuda -lz -ldl -lpthread
  rm -f "$OUT/tcnn_caller.o"
  echo "integration/build.sh: $OUT/tcnn_caller_ref" >&2
fi
