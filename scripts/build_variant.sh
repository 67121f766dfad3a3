#!/bin/bash
# build_variant.sh NAME [-Dflags...]: liborigami_b200 with extra defines -> gpurun_out/../variants/NAME.so (development aid)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --fmad=false -Xcompiler -fPIC -Xcompiler -O2 -shared "$@" \
  origami_b200/csrc/og_api.cu origami_b200/csrc/og_page.cu origami_b200/csrc/og_lines.cu origami_b200/csrc/og_strips.cu origami_b200/csrc/og_binarize.cu origami_b200/csrc/og_sauvola_warp.cu origami_b200/csrc/og_segment.cu origami_b200/csrc/og_handoff.cu origami_b200/csrc/og_geometry.cu \
  -o variants/$name.so
