#!/bin/bash
# build_variant.sh <name> <extra nvcc flags...>: builds vspg-pbrt-v4_b200/libvspg_<name>.so (same ABI) for A/B runs
set -e
cd "$(dirname "$0")/../vspg-pbrt-v4_b200/csrc"
name=$1; shift
mkdir -p build/$name
NV="/usr/local/cuda/bin/nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC,-O2,-ffp-contract=off -Xptxas -v --expt-relaxed-constexpr"
$NV "$@" -c vspg_ctx.cu -o build/$name/vspg_ctx.o 2> build/$name/ptxas.log || (cat build/$name/ptxas.log; false)
make -s build/host_exr.o build/host_field.o build/host_scene.o build/host_nvdb.o build/hg_table_data.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libvspg_$name.so build/$name/vspg_ctx.o build/host_exr.o build/host_field.o build/host_scene.o build/host_nvdb.o build/hg_table_data.o -lcudart_static -lz -lpthread -ldl -lrt
echo built libvspg_$name.so
