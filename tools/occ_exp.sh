#!/bin/bash
# needs the A/B build first (not made by the Makefile):
#   cd baitstr_b200/csrc && nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC \
#     -DBSTR_MAXTHREADS=224 -DBSTR_MINCTAS=3 -c select_host.cu -o /tmp/occ3.o && nvcc -shared -gencode arch=compute_100a,code=sm_100a \
#     -o ../libbaitstr_select_occ3.so /tmp/occ3.o stream_host.o pinflate.o -lz -lpthread -cudart shared
# occupancy experiment: phases A-E only (BSTR_EXP 4|8 = no format, no flush; +16 = dummy stage) at 2 / 3 CTAs per SM
run() { # label, lib, exp, geom
  BSTR_LIB=$2 BSTR_EXP=$3 BSTR_GEOM=$4 python bench.py --reads 4000000 --steps 3 --warmup 3 --quick 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', 'lib=$(basename $2)', 'exp=$3 geom=$4', 'ms/step', d['ms_per_step'], 'reads/s %.4g' % d['value'], 'kernel_us', d['roofline']['kernel_us_per_launch'])"
}
B=$PWD/baitstr_b200/libbaitstr_select.so
O=$PWD/baitstr_b200/libbaitstr_select_occ3.so
run full-base   $B 0  ""
run full-occ3lib $O 0  ""
run AE-2cta     $B 12 ""
run AE-2cta-dummy $B 28 "40960,224,4096"
run AE-3cta     $O 28 "40960,224,4096"
run AE-3cta-32k $O 28 "32768,192,4096"
run AE-4cta-24k $O 28 "24576,160,4096"
