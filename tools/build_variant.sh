#!/bin/bash
# build_variant.sh NAME [nvcc -D flags...]: compiles csrc/attention_tc.cu with the extra flags and links it with the
# objects of the current default build into minddiffusion_b200/libmdx_NAME.so (for A/B runs through MDX_LIB).
set -e
PKG=/root/repo/minddiffusion_b200
name=$1; shift
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -DMDX_PRECISE \
  -I/root/repo/include "$@" -c $PKG/csrc/attention_tc.cu -o /tmp/att_$name.o 2>&1 | grep -E "error|spill" || true
/usr/local/cuda/bin/nvcc -shared -o $PKG/libmdx_$name.so $(ls $PKG/build/*.o | grep -v attention_tc.o) /tmp/att_$name.o \
  -gencode arch=compute_100a,code=sm_100a -cudart static
ls -la $PKG/libmdx_$name.so
