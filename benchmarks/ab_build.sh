#!/bin/bash
# A/B builds of the persistent decode kernel: benchmarks/ab_build.sh <name> [-DFLAG=V ...]  ->  gpurun_ab/lib_<name>.so
# (select at run time with AO_B200_LIB=<path>; developer switch in _lib.py)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p gpurun_ab
python -m torchao_fork_b200.build >/dev/null
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
  --expt-relaxed-constexpr -DAO_B200_BUILD "$@" -c torchao_fork_b200/csrc/decode_mega.cu -o gpurun_ab/decode_mega_$name.o
objs=$(ls torchao_fork_b200/build/*.o | grep -v decode_mega.o)
nvcc -shared -o gpurun_ab/lib_$name.so $objs gpurun_ab/decode_mega_$name.o -gencode arch=compute_100a,code=sm_100a -cudart static
rm gpurun_ab/decode_mega_$name.o
echo gpurun_ab/lib_$name.so
