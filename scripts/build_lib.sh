#!/bin/bash
# Build libtile_engine.so in-tree for sm_100a.
set -e
cd "$(dirname "$0")/.."
mkdir -p slate_d35_b200/lib
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Iinclude \
     -shared -Xcompiler -fPIC "$@" -o slate_d35_b200/lib/libtile_engine.so slate_d35_b200/csrc/*.cu
