#!/bin/bash
# One 8-GPU session (gpurun --gpus 8): the >= 2-GPU tests with nothing skipped, the 8-GPU bench line with its
# parity self-check (both exchanges), and the config-5 deviation run (N = 1,048,576, 1000 steps, fast and FP64-force
# trajectories, FP64 energy every 100 steps, records through the reference's CheckpointManager).  Logs -> gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2_8gpu_devices.log 2>&1
timeout 1200 python -m pytest -m gpu -rs -q tests/test_gpu_multi.py tests/test_gpu_cpp_host.py tests/test_gpu_headless_app.py > gpurun_out/r2_multi_gpu_pytest_8gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_multi_gpu_pytest_8gpu.log
tail -5 gpurun_out/r2_multi_gpu_pytest_8gpu.log
timeout 400 $TR --master-port 29701 bench.py --gpus 8 --steps 6 --warmup 3 > gpurun_out/r2_bench_8gpu_p2p.log 2>&1; echo "bench p2p rc=$?"
timeout 500 $TR --master-port 29703 tools/run_config5.py --particles 1048576 --steps 1000 --every 100 --mode fast --out gpurun_out/r2_config5dev_fast.json --traj /tmp/c5_fast.bin > gpurun_out/r2_config5dev_fast.log 2>&1; echo "config5 fast rc=$?"
timeout 700 $TR --master-port 29704 tools/run_config5.py --particles 1048576 --steps 1000 --every 100 --mode fp64 --out gpurun_out/r2_config5dev_fp64.json --traj /tmp/c5_fp64.bin > gpurun_out/r2_config5dev_fp64.log 2>&1; echo "config5 fp64 rc=$?"
tail -3 gpurun_out/r2_config5dev_fast.log gpurun_out/r2_config5dev_fp64.log
