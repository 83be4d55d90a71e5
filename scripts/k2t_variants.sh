#!/bin/bash
# Tuning builds of K2t made HERE:  scripts/k2t_variants.sh tag1="-DX=1" ...  -> simplefe_b200/build/variants/libsfegpu_<tag>.so
set -e
cd "$(dirname "$0")/.."
python -m simplefe_b200.build > /dev/null
V=simplefe_b200/build/variants; mkdir -p $V
OTHERS=$(ls simplefe_b200/build/*.o | grep -v toeplitz_mma_kernel.cu.o)
for spec in "$@"; do
  tag=${spec%%=*}; flags=${spec#*=}
  ( nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr \
      $flags -c simplefe_b200/csrc/toeplitz_mma_kernel.cu -o $V/toep_$tag.o &&
    nvcc -shared -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -o $V/libsfegpu_$tag.so $OTHERS $V/toep_$tag.o && echo "built $tag ($flags)" ) &
done
wait
