#!/bin/bash
# A/B builds of the round kernel: tools/build_variant_lib.sh <name> <nvcc flags...> compiles csrc/irls_round.cu
# with the flags, links it with the regular objects of the other sources into
# ate_replication_causalml_b200/libate_b200_<name>.so; load it with ATE_B200_LIB=<path>.
set -e
name=$1; shift
cd "$(dirname "$0")/../ate_replication_causalml_b200"
mkdir -p build_var
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr "$@" -c csrc/irls_round.cu -o build_var/irls_round_$name.o
objs=$(ls build/*.o | grep -v irls_round.o)
/usr/local/cuda/bin/nvcc -shared -o libate_b200_$name.so $objs build_var/irls_round_$name.o -gencode arch=compute_100a,code=sm_100a
echo built libate_b200_$name.so
