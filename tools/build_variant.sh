#!/bin/bash
# tools/build_variant.sh NAME FILE.cu "<-D defines>" — a tuning variant of libdxgpu.so built HERE (no GPU needed): FILE.cu is
# recompiled with the defines, every other object comes from the normal build.  Output: data-accelerator_b200/variants/libdxgpu_NAME.so
# (git-ignored, travels to the GPU box); select it with DXGPU_LIB=... when running bench.py there.
set -e
NAME=$1; FILE=$2; DEFS=$3
cd "$(dirname "$0")/../data-accelerator_b200/csrc"
make -s
mkdir -p ../variants build_var
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -I../../include --expt-relaxed-constexpr -diag-suppress 177"
OBJS=$(ls build/*.o); VOBJS=""
for F in $FILE; do      # FILE may name several sources ("a.cu b.cu"): all of them are rebuilt with the defines
    $NV $DEFS -c $F -o build_var/${NAME}_${F%.cu}.o &
    OBJS=$(echo "$OBJS" | grep -v "build/${F%.cu}.o"); VOBJS="$VOBJS build_var/${NAME}_${F%.cu}.o"
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libdxgpu_$NAME.so $OBJS $VOBJS -lcudart
cuobjdump -res-usage ../variants/libdxgpu_$NAME.so 2>/dev/null | grep -A1 "k1_decode" | tail -1
