#!/bin/bash
# Build libcstone_b200 variants that differ in -D flags of neighbors.cu:  tools/nb_variants.sh name1:"-DX=1" name2:"-DY=2" ...
# and (on the GPU box) time each with: tools/nb_variants.sh --run name1 name2 ...
set -e
cd "$(dirname "$0")/.."
CS=cornerstone_octree_b200/csrc
if [ "$1" == "--run" ]; then
    shift
    cp cornerstone_octree_b200/libcstone_b200.so /tmp/lib_default.so
    for v in "$@"; do
        cp build/variants/lib_$v.so cornerstone_octree_b200/libcstone_b200.so
        python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e ${NB_BENCH_ARGS} 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v', 'ms/step', round(j['ms_per_step'],2), 'find_neighbors', j['stages']['find_neighbors']['ms'])"
    done
    cp /tmp/lib_default.so cornerstone_octree_b200/libcstone_b200.so
    exit 0
fi
mkdir -p build/variants
make -C $CS >/dev/null
for spec in "$@"; do
    name=${spec%%:*}; flags=${spec#*:}
    /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC $flags -c $CS/neighbors.cu -o build/variants/nb_$name.o
    objs=$(ls $CS/build/*.o | grep -v neighbors.o)
    /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/variants/lib_$name.so $objs build/variants/nb_$name.o -lnccl
    echo built $name "($flags)"
done
