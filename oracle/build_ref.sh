#!/bin/bash
# Builds oracle/_ref/libref_host.so: the reference's own header functions compiled for the host
# (see oracle/ref_host.cu). Needs /root/reference (the authoring container); the GPU box uses the
# prebuilt file if it travelled, and the committed golden vectors otherwise. Outputs only into oracle/_ref/.
set -eu
HERE="$(cd "$(dirname "$0")" && pwd)"
R="${NGP_REFERENCE:-/root/reference}"
[ -d "$R/include/neural-graphics-primitives" ] || { echo "[build_ref] $R not present: skipped"; exit 0; }
T="$R/dependencies/tiny-cuda-nn"
mkdir -p "$HERE/_ref"
nvcc -std=c++17 --extended-lambda --expt-relaxed-constexpr -O1 -w -shared -Xcompiler -fPIC -Xcompiler -ffp-contract=off \
	-I"$R/include" -I"$R/dependencies" -I"$T/include" -I"$T/dependencies" -I"$T/dependencies/fmt/include" \
	-DTCNN_MIN_GPU_ARCH=100 -DFMT_HEADER_ONLY -gencode arch=compute_100a,code=sm_100a \
	"$HERE/ref_host.cu" -o "$HERE/_ref/libref_host.so"
echo "[build_ref] built $HERE/_ref/libref_host.so"
g++ -std=c++17 -O1 -w -shared -fPIC -I"$T/dependencies" -I"$R/dependencies/NaturalSort" "$HERE/ref_json.cc" -o "$HERE/_ref/libref_json.so"
echo "[build_ref] built $HERE/_ref/libref_json.so"
g++ -std=c++17 -O1 -w -shared -fPIC -I"$R/dependencies/stb_image" -I"$R/dependencies/tinyexr" "$HERE/ref_images.cc" -o "$HERE/_ref/libref_images.so"
echo "[build_ref] built $HERE/_ref/libref_images.so"
