#!/bin/bash
# tile-geometry sweep of the fused kernel: BSTR_GEOM=tile,threads (0 = automatic threads)
# usage: tools/geom_sweep.sh WORKLOAD "tile,threads ..."
W=${1:-C2}; shift
for g in ${@:-"0,0"}; do
  BSTR_GEOM=$g python bench.py --workload $W --reads ${READS:-4000000} --steps 3 --warmup 3 --quick 2>&1 | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$W geom=$g', 'ms/step', d['ms_per_step'], 'Greads/s %.3f' % (d['value']/1e9), 'pipe_frac', d['roofline']['pipeline_frac_of_hbm'], 'sel', d['config']['selected_fraction'])
except Exception as e: print('$W geom=$g FAILED', e)"
done
