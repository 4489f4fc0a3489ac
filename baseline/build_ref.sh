#!/usr/bin/env bash
# Builds the UNMODIFIED reference (instant-ngp + its vendored tiny-cuda-nn) headless, from the sources
# where they lie under /root/reference, into the git-ignored baseline/_ref/ (which travels to the GPU
# box like our own .so).  Nothing of the reference is copied into the repository.
#
# This is a direct nvcc recipe, not the reference's CMake project: the three CMake options that matter
# are restated as flags (CMakeLists.txt:24-27,80-82: GUI / Vulkan / OptiX off, --use_fast_math,
# --extended-lambda, --expt-relaxed-constexpr; separable compilation :331), the glfw existence check
# (CMakeLists.txt:41-47; dependencies/glfw is empty in this snapshot) does not arise, and tcnn's arch
# cap at 90 (dependencies/tiny-cuda-nn/CMakeLists.txt:143-175) is lifted: native sm_100 SASS, so the
# baseline is the strongest build of the reference's own code this toolchain can make.
#   ARCH=100 (default)   native SASS for B200
#   ARCH=90              the as-shipped cap: compute_90 PTX only, JIT-compiled by the driver at load
# Outputs:
#   baseline/_ref/obj/libtcnn_ref.a        tiny-cuda-nn (10 TUs + fmt)
#   baseline/_ref/obj/libngp_ref.a         instant-ngp headless (16 TUs), device-linked
#   baseline/_ref/ref_harness          baseline/harness.cu linked against both
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=${OUT:-$HERE/_ref}
ARCH=${ARCH:-100}
JOBS=${JOBS:-$(nproc)}
TCNN=$REF/dependencies/tiny-cuda-nn
mkdir -p "$OUT/obj"

if [ ! -d "$REF/src" ]; then
  echo "build_ref: $REF absent (GPU box): using the prebuilt files in $OUT" >&2
  exit 0
fi

if [ "$ARCH" = "90" ]; then
  GENCODE="-gencode arch=compute_90,code=compute_90"
else
  GENCODE="-gencode arch=compute_${ARCH},code=sm_${ARCH}"
fi

cat > "$OUT/Makefile" <<EOF
REF := $REF
TCNN := $TCNN
OUT := $OUT
NVCC := nvcc
GENCODE := $GENCODE
COMMON := -std=c++14 -O3 -DNDEBUG --extended-lambda --expt-relaxed-constexpr \\
  -Xcompiler=-fPIC,-fno-strict-aliasing,-Wno-float-conversion \\
  -Xcudafe=--diag_suppress=unrecognized_gcc_pragma -w \\
  -DTCNN_MIN_GPU_ARCH=$ARCH \$(GENCODE)
TCNN_INC := -I\$(TCNN)/include -I\$(TCNN)/dependencies -I\$(TCNN)/dependencies/cutlass/include \\
  -I\$(TCNN)/dependencies/cutlass/tools/util/include -I\$(TCNN)/dependencies/fmt/include
NGP_INC := -I\$(REF)/include -I\$(REF)/dependencies -I\$(REF)/dependencies/filesystem \\
  -I\$(REF)/dependencies/nanovdb -I\$(REF)/dependencies/NaturalSort -I\$(REF)/dependencies/tinylogger \\
  -I\$(REF)/dependencies/zstr/src
NGP_FLAGS := --use_fast_math -rdc=true -DNGP_VERSION='"1.0dev"'

TCNN_SRCS := common_host cpp_api cutlass_mlp encoding loss network object optimizer reduce_sum fully_fused_mlp
NGP_CU := camera_path common_host marching_cubes nerf_loader render_buffer testbed testbed_image \\
  testbed_nerf testbed_sdf testbed_geometry testbed_volume tinyexr_wrapper tinyobj_loader_wrapper \\
  triangle_bvh geometry_bvh
TCNN_OBJS := \$(addprefix \$(OUT)/obj/tcnn_,\$(addsuffix .o,\$(TCNN_SRCS))) \$(OUT)/obj/fmt_format.o \$(OUT)/obj/fmt_os.o
NGP_OBJS := \$(addprefix \$(OUT)/obj/ngp_,\$(addsuffix .o,\$(NGP_CU))) \$(OUT)/obj/ngp_thread_pool.o

all: \$(OUT)/obj/libtcnn_ref.a \$(OUT)/obj/libngp_ref.a

# long poles first so make -j schedules them early
\$(OUT)/obj/tcnn_%.o: \$(TCNN)/src/%.cu
	\$(NVCC) \$(COMMON) \$(TCNN_INC) -c \$< -o \$@
\$(OUT)/obj/fmt_%.o: \$(TCNN)/dependencies/fmt/src/%.cc
	g++ -std=c++14 -O2 -fPIC -I\$(TCNN)/dependencies/fmt/include -c \$< -o \$@
\$(OUT)/obj/ngp_%.o: \$(REF)/src/%.cu
	\$(NVCC) \$(COMMON) \$(NGP_FLAGS) \$(TCNN_INC) \$(NGP_INC) -c \$< -o \$@
\$(OUT)/obj/ngp_thread_pool.o: \$(REF)/src/thread_pool.cpp
	g++ -std=c++14 -O2 -fPIC -DTCNN_MIN_GPU_ARCH=$ARCH -I\$(REF)/include -I\$(REF)/dependencies \$(TCNN_INC) -I/usr/local/cuda/include -c \$< -o \$@

\$(OUT)/obj/libtcnn_ref.a: \$(TCNN_OBJS)
	rm -f \$@ && ar rcs \$@ \$^
\$(OUT)/obj/ngp_dlink.o: \$(NGP_OBJS)
	\$(NVCC) \$(GENCODE) -Xcompiler=-fPIC -dlink \$(filter %.o,\$^) -o \$@
\$(OUT)/obj/libngp_ref.a: \$(NGP_OBJS) \$(OUT)/obj/ngp_dlink.o
	rm -f \$@ && ar rcs \$@ \$^
EOF

echo "build_ref: ARCH=$ARCH JOBS=$JOBS -> $OUT" >&2
# the two long translation units first, then everything else
make -C "$OUT" -f "$OUT/Makefile" -j"$JOBS" -k all 2>&1 | tee "$OUT/build.log" | grep -v "^make" || true
ls -la "$OUT"/obj/*.a >&2 || { echo "build_ref: library build failed, see $OUT/build.log" >&2; exit 1; }

if [ -f "$HERE/harness.cu" ]; then
  "$HERE/build_harness.sh"
fi
