#!/bin/bash
# build_variant.sh NAME "EXTRA nvcc flags"  ->  gpurun_out is scratch; variants live in fl-rl-compression-mpi_b200/variants/ (git-ignored .so)
set -e
HERE=$(cd "$(dirname "$0")/.." && pwd)
PKG=$HERE/fl-rl-compression-mpi_b200
NAME=$1; shift
mkdir -p $PKG/variants/build_$NAME
for f in $PKG/csrc/*.cu; do
  o=$PKG/variants/build_$NAME/$(basename ${f%.cu}).o
  nvcc -std=c++17 -O3 "$@" -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -I$HERE/include -c $f -o $o &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $PKG/variants/libflrl_$NAME.so $PKG/variants/build_$NAME/*.o -cudart static -ldl -lpthread
rm -rf $PKG/variants/build_$NAME
echo built $PKG/variants/libflrl_$NAME.so
