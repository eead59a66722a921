set -x
nvidia-smi -L | wc -l; nproc; free -g | head -2
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_gpu" > gpurun_out/r02_pytest_8gpu.log 2>&1; tail -3 gpurun_out/r02_pytest_8gpu.log
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02_bench_8gpu.json 2> gpurun_out/r02_bench_8gpu.err ) 2>&1 | tail -3; echo rc=$?; tail -c 800 gpurun_out/r02_bench_8gpu.err
