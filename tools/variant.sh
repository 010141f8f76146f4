#!/bin/bash
# Build an experimental variant of libgnncca.so next to the product library:
#   tools/variant.sh NAME -DFLAG1 -DFLAG2=3 ...      ->  gnn-cca_b200/variants/libgnncca_NAME.so
# Run anything with it:  GNNCCA_LIB=gnn-cca_b200/variants/libgnncca_NAME.so python tools/stepbench.py
set -e
cd "$(dirname "$0")/.."
name=$1; shift
out=gnn-cca_b200/variants; mkdir -p $out/obj_$name
for f in gnn-cca_b200/csrc/*.cu; do
  o=$out/obj_$name/$(basename ${f%.cu}).o
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr "$@" -c $f -o $o &
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $out/libgnncca_$name.so $out/obj_$name/*.o -lcudart
echo built $out/libgnncca_$name.so
