#!/bin/bash
# usage: scripts/build_variant.sh <name> <-D flags for direct.cu ...>: wormhole_b200/variants/lib_<name>.so = the current
# objects with direct.cu recompiled under the given macros (WH_DX_MAXNREG, WH_DX_GATHER, WH_DX_ILP). Select with WH_LIB.
set -e
name=$1; shift
cd "$(dirname "$0")/.."
mkdir -p wormhole_b200/variants
python -c "from wormhole_b200 import build as b; b.build()"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC,-fvisibility=hidden,-O2 -ccbin /usr/bin/g++ -Iinclude "$@" -c wormhole_b200/csrc/direct.cu -o wormhole_b200/variants/direct_$name.o
objs=$(ls wormhole_b200/csrc/build/*.o | grep -v "/direct.o")
nvcc -shared -o wormhole_b200/variants/lib_$name.so $objs wormhole_b200/variants/direct_$name.o -lcudart -ccbin /usr/bin/g++
rm wormhole_b200/variants/direct_$name.o
echo wormhole_b200/variants/lib_$name.so
