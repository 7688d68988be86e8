#!/bin/bash
# A/B measurement build of libdrt.so with extra -D flags: tools/build_variant.sh <name> <flags...>  ->  dustraytracer_b200/ab/libdrt_<name>.so
# (used with DRT_LIB_PATH=... python tools/sweep.py ...; the directory is git-ignored and travels to the GPU box)
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
src=$root/dustraytracer_b200/csrc
out=$root/dustraytracer_b200/ab
obj=$out/obj_$name
mkdir -p "$obj"
FLAGS="-std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-fno-fast-math,-ffp-contract=off -ccbin /usr/bin/g++"
for f in drt_kernels.cu drt_api.cu drt_gltf.cpp drt_gltf_robust.cpp drt_png.cpp drt_jpeg.cpp; do
  nvcc $FLAGS "$@" -c "$src/$f" -o "$obj/$f.o" &
done
wait
nvcc -shared -ccbin /usr/bin/g++ -Xlinker -Bsymbolic -o "$out/libdrt_$name.so" "$obj"/*.o -lz
rm -rf "$obj"
ls -la "$out/libdrt_$name.so"
