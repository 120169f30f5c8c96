#!/bin/bash
# sort_bench.py on every library variant under gsplat-triton_b200/lib/variants (built with other GSB_SORT_* settings).
# A variant = radix_sort.cu recompiled with other tile shapes, linked with the objects of the regular build:
#   F="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --use_fast_math -Xcompiler -fPIC -I include -I gsplat-triton_b200/csrc"
#   nvcc $F -DGSB_SORT_ITEMS=12 -DGSB_SORT_MIN_BLOCKS=2 -c gsplat-triton_b200/csrc/radix_sort.cu -o /tmp/rs.o
#   nvcc -shared -gencode arch=compute_100a,code=sm_100a -o gsplat-triton_b200/lib/variants/sort_i12_b2.so \
#        $(ls gsplat-triton_b200/build/*.o | grep -v radix_sort.o) /tmp/rs.o
# (lib/ is git-ignored and travels with gpurun; GSPLAT_B200_LIB selects the library to load)
mkdir -p gpurun_out
echo "default:"; python tools/sort_bench.py 2>&1 | tail -3
for so in gsplat-triton_b200/lib/variants/sort_*.so; do
    echo "$(basename $so):"; GSPLAT_B200_LIB=$PWD/$so python tools/sort_bench.py 2>&1 | tail -3
done
