#!/bin/bash
# phase attribution: time the fused kernel with phases skipped (BSTR_EXP bits: 1 match, 2 validate, 4 format, 8 flush)
for e in ${EXPS:-0 1 3 4 8 12 15}; do
  BSTR_EXP=$e python bench.py --workload ${WL:-C2} --quick --reads ${READS:-4000000} --steps 3 --warmup 2 2>&1 | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('exp=$e', 'ms/step', d['ms_per_step'], 'reads/s %.3g' % d['value'], 'sel', d['config']['selected_fraction'])"
done
