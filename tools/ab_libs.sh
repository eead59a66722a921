#!/bin/bash
# A/B kernel builds: tools/ab_libs.sh <workload> lib1.so lib2.so ...   (kernel-only bench, 3 repeats each)
WL=$1; shift
for rep in 1 2 3; do for L in "$@"; do
  BSTR_LIB=$PWD/$L python bench.py --workload $WL --reads 4000000 --steps 5 --warmup 3 --quick 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$WL', '$(basename $L)', 'reads/s %.4g' % d['value'], 'kernel_us', d['roofline']['kernel_us_per_launch'], d['out_checksum_fnv1a'])"
done; done
