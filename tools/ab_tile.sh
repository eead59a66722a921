for w in C2 C3 C4 C5; do for a in 0 1; do BSTR_AUTOTILE=$a python bench.py --workload $w --quick 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w autotile=$a', 'Greads/s %.3f' % (d['value']/1e9), d['ms_per_step'], d['roofline']['frac'], d['out_checksum_fnv1a'])"; done; done
