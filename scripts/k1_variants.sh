#!/bin/bash
# Tuning builds of K1 made HERE (nvcc cross-compiles) so that an A/B on the GPU box costs no compile time:
#   scripts/k1_variants.sh tag1="-DX=1 -DY=0" tag2="..."      -> simplefe_b200/build/variants/libsfegpu_<tag>.so
# On the box: cp simplefe_b200/build/variants/libsfegpu_<tag>.so simplefe_b200/libsfegpu.so; python bench.py ...
set -e
cd "$(dirname "$0")/.."
python -m simplefe_b200.build > /dev/null
V=simplefe_b200/build/variants; mkdir -p $V
OTHERS=$(ls simplefe_b200/build/*.o | grep -v blkconv_kernel.cu.o)
for spec in "$@"; do
  tag=${spec%%=*}; flags=${spec#*=}
  ( nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr \
      $flags -c simplefe_b200/csrc/blkconv_kernel.cu -o $V/blkconv_$tag.o &&
    nvcc -shared -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -o $V/libsfegpu_$tag.so $OTHERS $V/blkconv_$tag.o && echo "built $tag ($flags)" ) &
done
wait
