#!/bin/bash
# Build libmfb_<name>.so with extra -D flags for the update kernel (MFB_V4_THREADS, MFB_V4_MINB, MFB_V4_ROLL) next to the
# default library, for A/B runs on the GPU box:  tools/build_variant.sh A -DMFB_V4_MINB=2 ; MFB_LIB=... python bench.py
set -e
name=$1; shift
cd "$(dirname "$0")/../mixed-fem_b200/csrc"
make -s >/dev/null
mkdir -p ../exp /tmp/mfb_var
NV=/usr/local/cuda/bin/nvcc
$NV -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr "$@" \
    -c mfb_api.cu -o /tmp/mfb_var/api_$name.o 2> /tmp/mfb_var/api_$name.log
$NV -gencode arch=compute_100a,code=sm_100a -shared -o ../exp/libmfb_$name.so mfb_create.o /tmp/mfb_var/api_$name.o mfb_fields.o \
    mfb_pcg.o mfb_newton.o mfb_dist.o mfb_saddle.o mfb_stretch.o mfb_partition.o mfb_noalloc.o -lcudart -ldl
grep -A2 "k_update_tet_v4ILb1ELi1E" /tmp/mfb_var/api_$name.log | tail -2
