set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_v11.log 2>&1; tail -2 gpurun_out/pytest_gpu_v11.log
python bench.py > gpurun_out/bench_v11.json 2> gpurun_out/bench_v11.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_v11.json 2> gpurun_out/bench_ref_v11.err
for w in C3 C4 C5; do python bench.py --workload $w --no-ingest > gpurun_out/bench_${w}_v11.json 2>/dev/null; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v11.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-ingest > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_fused_tile -s 3 -c 2 -f -o gpurun_out/fused_v11 python bench.py --reads 2000000 --steps 2 --warmup 1 --no-cpu --no-ingest > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out/
