#!/bin/bash
# Experiment builds of the CUDA engine with extra nvcc flags (A/B runs of compile-time knobs):
#   tools/build_variant.sh <name> [nvcc flags...]   ->  gpu-ray-tracer_b200/variants/librtb200_<name>.so
# run with RTB200_CUDA_LIB=gpu-ray-tracer_b200/variants/librtb200_<name>.so python bench.py ...
set -e
cd "$(dirname "$0")/.."
NAME=$1; shift
mkdir -p gpu-ray-tracer_b200/variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC -shared -Iinclude \
  -Igpu-ray-tracer_b200/csrc "$@" gpu-ray-tracer_b200/csrc/rt_engine.cu gpu-ray-tracer_b200/csrc/rt_bvh_build.cu \
  -o gpu-ray-tracer_b200/variants/librtb200_${NAME}.so
echo built gpu-ray-tracer_b200/variants/librtb200_${NAME}.so
