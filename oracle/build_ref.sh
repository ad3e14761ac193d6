#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY. Builds the UNMODIFIED reference CUDA sources, where
# they lie under /root/reference, into oracle/_ref/ (git-ignored, travels to the
# GPU box with gpurun). No reference source is copied into this repository.
#
#   oracle/_ref/libref_dgr.so  <- DGR/cuda_rasterizer/{forward,backward,rasterizer_impl}.cu
#                                 + oracle/ref_cuda/ref_capi.cu (our C-ABI shim)
#   oracle/_ref/libref_knn.so  <- KNN/simple_knn.cu + oracle/ref_cuda/ref_knn_capi.cu
#   oracle/_ref/libref_utils.so <- TS/utils/cuda_utils.cu + oracle/ref_cuda/ref_utils_capi.cu
#                                 (gaussian2occgrid, contour_padding; plain CUDA + thrust, no torch)
#
# Flags mirror the reference build (DGR/setup.py:18-29: only -I third_party/glm,
# no fast-math, default -fmad=true); the arch is the only addition.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${SOAR_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
DGR="$REF/submodules/diff-gaussian-rasterization"
KNN="$REF/submodules/simple-knn"
TSU="$REF/soar/threestudio-soar/utils"
ARCH="-gencode arch=compute_100a,code=sm_100a"
if [ ! -d "$DGR/cuda_rasterizer" ]; then
  echo "build_ref: $DGR not found (reference tree not mounted) — keeping prebuilt files" >&2
  exit 0
fi
mkdir -p "$OUT/obj"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
COMMON="-O3 -std=c++17 -Xcompiler -fPIC $ARCH -lineinfo -w"

build_obj() {  # src obj extra...
  local src="$1" obj="$2"; shift 2
  if [ ! -f "$obj" ] || [ "$src" -nt "$obj" ]; then
    $NVCC $COMMON "$@" -c "$src" -o "$obj"
  fi
}

build_obj "$DGR/cuda_rasterizer/forward.cu"         "$OUT/obj/forward.o"    -I"$DGR/third_party/glm" -I"$DGR/cuda_rasterizer" &
build_obj "$DGR/cuda_rasterizer/backward.cu"        "$OUT/obj/backward.o"   -I"$DGR/third_party/glm" -I"$DGR/cuda_rasterizer" &
build_obj "$DGR/cuda_rasterizer/rasterizer_impl.cu" "$OUT/obj/impl.o"       -I"$DGR/third_party/glm" -I"$DGR/cuda_rasterizer" &
build_obj "$HERE/ref_cuda/ref_capi.cu"              "$OUT/obj/ref_capi.o"   -I"$DGR/third_party/glm" -I"$DGR/cuda_rasterizer" &
build_obj "$KNN/simple_knn.cu"                      "$OUT/obj/simple_knn.o" -I"$KNN" &
build_obj "$HERE/ref_cuda/ref_knn_capi.cu"          "$OUT/obj/ref_knn_capi.o" -I"$KNN" &
build_obj "$TSU/cuda_utils.cu"                     "$OUT/obj/cuda_utils.o" -I"$TSU" &
build_obj "$HERE/ref_cuda/ref_utils_capi.cu"        "$OUT/obj/ref_utils_capi.o" -I"$TSU" &
wait

$NVCC -shared $ARCH -o "$OUT/libref_dgr.so" "$OUT/obj/forward.o" "$OUT/obj/backward.o" "$OUT/obj/impl.o" "$OUT/obj/ref_capi.o" -lcudart
$NVCC -shared $ARCH -o "$OUT/libref_knn.so" "$OUT/obj/simple_knn.o" "$OUT/obj/ref_knn_capi.o" -lcudart
$NVCC -shared $ARCH -o "$OUT/libref_utils.so" "$OUT/obj/cuda_utils.o" "$OUT/obj/ref_utils_capi.o" -lcudart
echo "build_ref: ok -> $OUT/libref_dgr.so $OUT/libref_knn.so $OUT/libref_utils.so"
