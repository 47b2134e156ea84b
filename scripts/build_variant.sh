#!/bin/bash
# Tuning aid: rebuild only centrality.cu (the one unit that instantiates the pull kernel) with
# extra -D flags and link it with the other objects into
# cugraph_forked_b200/variants/libcugraph_b200_<name>.so.  Select it with B200_LIBRARY=<path>.
#   scripts/build_variant.sh g24 -DB200_GATHER_WARPS=24
set -e
name=$1; shift
here=$(cd "$(dirname "$0")/.." && pwd)
cs=$here/cugraph_forked_b200/csrc
mkdir -p $cs/build/var_$name $here/cugraph_forked_b200/variants
make -C $cs -j8 > /dev/null
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a \
  -Xcompiler -fPIC --expt-relaxed-constexpr -I$here/include "$@" \
  -c $cs/centrality.cu -o $cs/build/var_$name/centrality.o
objs=$(ls $cs/build/*.o | grep -v centrality.o)
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a \
  -o $here/cugraph_forked_b200/variants/libcugraph_b200_$name.so $objs $cs/build/var_$name/centrality.o -lnccl
echo built $here/cugraph_forked_b200/variants/libcugraph_b200_$name.so
