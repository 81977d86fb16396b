#!/bin/bash
# Builds oracle/_ref/libnerf_ref.so: the REFERENCE's own kernels compiled for the host (test tooling, dev container
# only — /root/reference does not exist on the GPU box).  Sources stay under /root/reference; the two extracted
# header slices are generated into oracle/_ref/ (git-ignored), never committed.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
R="${NRF_REFERENCE_DIR:-/root/reference}"
OUT="$HERE/_ref"
[ -d "$R/include/nerf-cuda" ] || { echo "reference not present at $R — skipping oracle/_ref" >&2; exit 0; }
mkdir -p "$OUT"
T="$R/dependencies/tiny-cuda-nn"
{
  echo '// GENERATED at build time from tcnn encodings/grid.h:46-311 — do not commit'
  echo '#include <tiny-cuda-nn/common.h>'
  echo '#include <tiny-cuda-nn/common_device.h>'
  echo '#include <tiny-cuda-nn/encoding.h>'
  echo '#include <tiny-cuda-nn/random.h>'
  sed -n '46,311p' "$T/include/tiny-cuda-nn/encodings/grid.h"
  echo 'TCNN_NAMESPACE_END'
} > "$OUT/grid_extract.h"
{
  echo '// GENERATED at build time from tcnn encodings/spherical_harmonics.h:46-152 — do not commit'
  echo '#include <tiny-cuda-nn/common.h>'
  echo '#include <tiny-cuda-nn/gpu_matrix.h>'
  echo 'TCNN_NAMESPACE_BEGIN'
  sed -n '46,152p' "$T/include/tiny-cuda-nn/encodings/spherical_harmonics.h"
  echo 'TCNN_NAMESPACE_END'
} > "$OUT/sh_extract.h"
g++ -std=c++17 -O2 -ffp-contract=off -fopenmp -w -shared -fPIC \
  -I"$HERE/ref_harness/override" -I"$OUT" -I"$R/include" -I"$R/dependencies" -I"$R/dependencies/eigen" \
  -I"$R/dependencies/filesystem" -I"$R/dependencies/tinylogger" -I"$T/include" -I"$T/dependencies" \
  -I"$T/dependencies/fmt/include" -I/usr/local/cuda/include -DTCNN_MIN_GPU_ARCH=100 -DFMT_HEADER_ONLY \
  "$HERE/ref_harness/ref_harness.cpp" -o "$OUT/libnerf_ref.so" \
  -L/usr/local/cuda/lib64 -L/usr/local/cuda/lib64/stubs -lcudart -lcuda
echo "built $OUT/libnerf_ref.so"
