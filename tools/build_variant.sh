#!/bin/bash
# Build-container helper: a variant of libhugs_b200.so with ONE source file patched by a sed expression.
# usage: bash tools/build_variant.sh NAME file.cu 'sed-expression'   -> ml-hugs_b200/lib/variants/libhugs_b200_NAME.so
set -e
NAME=$1; SRC=$2; EXPR=$3
C=ml-hugs_b200/csrc; B=ml-hugs_b200/build; V=ml-hugs_b200/lib/variants
mkdir -p $V /tmp/hbvar_$NAME
sed "$EXPR" $C/$SRC > $C/_variant_$NAME.cu
if cmp -s $C/$SRC $C/_variant_$NAME.cu; then echo "sed expression changed nothing"; rm $C/_variant_$NAME.cu; exit 1; fi
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
  -Xptxas -v -c $C/_variant_$NAME.cu -o /tmp/hbvar_$NAME/v.o 2>&1 | grep -B1 -A1 "spill" | grep -v "^--" | grep -i "kernel\|spill" | sed 's/ptxas info    : //' | cut -c1-200 | grep -B1 "[1-9][0-9]* bytes spill" || true
rm $C/_variant_$NAME.cu
OBJS=$(ls $B/*.o | grep -v "/${SRC%.cu}.o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o $V/libhugs_b200_$NAME.so $OBJS /tmp/hbvar_$NAME/v.o
echo "built $V/libhugs_b200_$NAME.so"
