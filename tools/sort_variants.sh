#!/bin/bash
# Build libcstone_b200 variants that differ in -D flags of radix_sort.cu (sort probe):  tools/nb_variants.sh name1:"-DX=1" name2:"-DY=2" ...
# and (on the GPU box) time each with: tools/nb_variants.sh --run name1 name2 ...
set -e
cd "$(dirname "$0")/.."
CS=cornerstone_octree_b200/csrc
if [ "$1" == "--run" ]; then
    shift
    cp cornerstone_octree_b200/libcstone_b200.so /tmp/lib_default.so
    for v in "$@"; do
        cp build/variants/lib_$v.so cornerstone_octree_b200/libcstone_b200.so
        python tools/sort_probe.py 32000000 7 2>&1 | grep -v nearly | sed "s/^/$v /"
    done
    cp /tmp/lib_default.so cornerstone_octree_b200/libcstone_b200.so
    exit 0
fi
mkdir -p build/variants
make -C $CS >/dev/null
for spec in "$@"; do
    name=${spec%%:*}; flags=${spec#*:}
    /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC $flags -c $CS/radix_sort.cu -o build/variants/rs_$name.o
    objs=$(ls $CS/build/*.o | grep -v radix_sort.o)
    /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/variants/lib_$name.so $objs build/variants/rs_$name.o -lnccl
    echo built $name "($flags)"
done
