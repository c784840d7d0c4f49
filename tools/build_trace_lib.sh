#!/bin/bash
# A second build of the library with the step kernel's slot trace compiled in (-DATE_ROUND_TRACE);
# load it with ATE_B200_LIB=ate_replication_causalml_b200/libate_b200_trace.so.
set -e
cd "$(dirname "$0")/../ate_replication_causalml_b200"
mkdir -p build_trace
for f in csrc/*.cu; do
  o=build_trace/$(basename "${f%.cu}").o
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr -DATE_ROUND_TRACE -c "$f" -o "$o" &
done
wait
/usr/local/cuda/bin/nvcc -shared -o libate_b200_trace.so build_trace/*.o -gencode arch=compute_100a,code=sm_100a
echo built libate_b200_trace.so
