#!/bin/bash
# usage: tools/build_variant.sh <name> <extra nvcc flags...>  ->  kubernetes-machine-learning-server_b200/variants/libkmls_<name>.so
name=$1; shift
src=kubernetes-machine-learning-server_b200/csrc
obj=/tmp/kmls_variant_$name; mkdir -p $obj kubernetes-machine-learning-server_b200/variants
for f in errors level1 pairs pair_gemm levelk project rules mine rules_host exchange; do
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Iinclude -I$src \
    --expt-relaxed-constexpr -DCCCL_IGNORE_DEPRECATED_API -Xptxas -v "$@" -c $src/$f.cu -o $obj/$f.o 2> $obj/$f.log &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o kubernetes-machine-learning-server_b200/variants/libkmls_$name.so $obj/*.o -lcudart
grep -A3 "k_count_andE\|k_count_projectE" $obj/mine.log | grep -E "Used|spill"
