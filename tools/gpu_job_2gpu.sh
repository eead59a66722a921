# 2-GPU box: the >=2-GPU parity test of the in-process multi-GPU path, then bench.py under torchrun
set -x
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_gpu or run_files or async_lanes" > gpurun_out/r02_pytest_2gpu.log 2>&1; tail -3 gpurun_out/r02_pytest_2gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02_bench_2gpu.json 2> gpurun_out/r02_bench_2gpu.err; echo rc=$?; tail -c 600 gpurun_out/r02_bench_2gpu.err; nproc
