#!/bin/bash
# build an experimental fp32 product library with extra -D flags: scripts/build_variant.sh <name> <flags...>
set -e
name=$1; shift
cd "$(dirname "$0")/.."
out=variants/$name; mkdir -p $out/obj
for s in engine slab; do
  nvcc -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-O3,-fno-gnu-unique -DFFTW_SINGLE=1 "$@" -Iinclude -Ifftw3_b200/csrc -c fftw3_b200/csrc/$s.cu -o $out/obj/$s.o &
done
for s in fftw_api fftw_mpi_api; do
  gcc -O2 -fPIC -std=c99 -DFFTW_SINGLE=1 -Iinclude -Ifftw3_b200/csrc -c fftw3_b200/csrc/$s.c -o $out/obj/$s.o
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $out/libfftw3.so $out/obj/*.o -Xlinker -Bsymbolic -lcudart_static -ldl -lpthread -lrt
echo built $out/libfftw3.so
