# Final round-2 evidence on ONE GPU: the whole GPU suite, the bench line, the reference arm, one ncu capture of the extension kernel
set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > gpurun_out/r02_pytest_gpu_1gpu.log; cat gpurun_out/r02_pytest_gpu_1gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err || { tail -5 gpurun_out/r02_bench_1gpu.err; exit 1; }
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_ref.err
python tools/extend_bench.py 250000 250 30 150 > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_extend_blocks -s 1 -c 1 -f -o gpurun_out/r02_extend_blocks python tools/extend_bench.py 250000 250 30 150 > gpurun_out/ncu_e.log 2>&1
python -c "
import json
d=json.load(open('gpurun_out/r02_bench_1gpu.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], json.dumps(d['next_stages']), json.dumps(d['ingest']))
r=[json.loads(l) for l in open('gpurun_out/r02_bench_reference_arm.json') if l.startswith('{')][-1]; print('reference arm', r['value'], r.get('cpu_baseline'))"
