#!/bin/bash
# compile the library into gpurun_out/wip (not shipped by gpurun): register counts before GPU time is spent
# usage: tools/wipbuild.sh [name] [extra nvcc flags]  -> gpurun_out/wip/lib_<name>.so
set -e
cd /root/repo/baitstr_b200/csrc
N=${1:-wip}; shift || true
W=/root/repo/gpurun_out/wip; mkdir -p $W
ARCH="-gencode arch=compute_100a,code=sm_100a"
FL="-O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC,-Wall,-Wno-unused-function -Xptxas -v"
/usr/local/cuda/bin/nvcc $FL "$@" -c select_host.cu -o $W/select_host_$N.o 2> $W/ptxas_$N.log || { grep -i error $W/ptxas_$N.log | head; exit 1; }
[ -f stream_host.o ] || make stream_host.o pinflate.o >/dev/null
/usr/local/cuda/bin/nvcc -shared $ARCH -o $W/lib_$N.so $W/select_host_$N.o stream_host.o pinflate.o -lz -lpthread -cudart shared
grep -A2 "Compiling entry function '_ZN4bstr12k_fused_tile" $W/ptxas_$N.log | grep "Used\|spill" 
