#!/bin/bash
# build a tuning variant of libb200denoise.so:  scripts/build_variant.sh NAME "-DB2D_FT_TH=64 ..."
set -e
cd "$(dirname "$0")/../image_denoising_b200/csrc"
mkdir -p ../../variants/obj_$1
for f in *.cu; do
  o=../../variants/obj_$1/${f%.cu}.o
  if [ "$f" = "b2d_fused3w.cu" ] || [ "$f" = "b2d_march3.cu" ] || [ ! -f build/${f%.cu}.o ] || [ -n "$3" ]; then
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC $2 -c $f -o $o &
  else
    cp build/${f%.cu}.o $o
  fi
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../variants/lib_$1.so ../../variants/obj_$1/*.o -lcudart
rm -rf ../../variants/obj_$1
echo built variants/lib_$1.so
