#!/bin/bash
# warp-tile sweep of k_fused_warp: BSTR_WRECS = records per warp tile
# usage: tools/wgeom_sweep.sh WORKLOAD "recs ..."
W=${1:-C2}; shift
for r in ${@:-30.5}; do
  BSTR_WRECS=$r python bench.py --workload $W --reads ${READS:-4000000} --steps 3 --warmup 2 --no-cpu 2>&1 | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$W wrecs=$r', 'ms/step', d['ms_per_step'], 'Greads/s %.3f' % (d['value']/1e9), 'pipe_frac', d['roofline']['pipeline_frac_of_hbm'], 'sel', d['config']['selected_fraction'])
except Exception as e: print('$W wrecs=$r FAILED', e)"
done
