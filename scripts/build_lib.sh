#!/bin/bash
# Build libtile_engine.so in-tree for sm_100a.
set -e
cd "$(dirname "$0")/.."
mkdir -p slate_d35_b200/lib
NCCL_LIB=$(python -c "import os, nvidia.nccl as n; print(os.path.join(list(n.__path__)[0], 'lib'))" 2>/dev/null || echo /usr/lib/x86_64-linux-gnu)
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Iinclude \
     -shared -Xcompiler -fPIC "$@" -o slate_d35_b200/lib/libtile_engine.so slate_d35_b200/csrc/*.cu \
     -L"$NCCL_LIB" -L/usr/lib/x86_64-linux-gnu -l:libnccl.so.2 \
     -Xlinker -rpath -Xlinker "$NCCL_LIB" -Xlinker -rpath -Xlinker /usr/lib/x86_64-linux-gnu
