#!/bin/bash
# A/B builds of one kernel file with extra -D flags: tools/build_variant.sh NAME k_file.cu "-DFOO=1 -DBAR=2"
# -> gf-orb-slam2_b200/variants/libgforb_NAME.so (git-ignored; select with GFORB_LIBRARY=...)
set -e
cd "$(dirname "$0")/../gf-orb-slam2_b200/csrc"
mkdir -p ../variants _build/var
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden \
  --fmad=false -cudart static $3 -c $2 -o _build/var/$1.o
OBJS=$(ls _build/*.o | grep -v "_build/${2%.cu}.o")
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o ../variants/libgforb_$1.so $OBJS _build/var/$1.o
