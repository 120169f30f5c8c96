#!/bin/bash
# sort_bench.py on every library variant under gsplat-triton_b200/lib/variants (built with other GSB_SORT_* settings)
mkdir -p gpurun_out
echo "default:"; python tools/sort_bench.py 2>&1 | tail -3
for so in gsplat-triton_b200/lib/variants/sort_*.so; do
    echo "$(basename $so):"; GSPLAT_B200_LIB=$PWD/$so python tools/sort_bench.py 2>&1 | tail -3
done
