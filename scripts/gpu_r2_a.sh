#!/bin/bash
# round 2, call A: parity tests incl. the termination rounds, the reference-CUDA goldens + timings, the L2 probe,
# and a short C3 bench (no regression from the round refactor).  Every step logs to gpurun_out/ and never aborts the rest.
mkdir -p gpurun_out/ref_cuda
nvidia-smi -L > gpurun_out/r2a_gpu.txt 2>&1
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40) > gpurun_out/r2a_pytest.log
(timeout 300 python baseline/run_ref_cuda.py golden --out gpurun_out/ref_cuda 2>&1 | tail -20) > gpurun_out/r2a_ref_golden.log
(timeout 600 python baseline/run_ref_cuda.py time --workloads C2,C3 --reps 3 --out gpurun_out/ref_cuda 2>&1 | tail -20) > gpurun_out/r2a_ref_time.log
(timeout 200 python scripts/mem_probe.py 2>&1 | tail -20) > gpurun_out/r2a_mem_probe.log
(timeout 300 python bench.py --steps 16 --warmup 4 --no-cpu 2>&1 | tail -5) > gpurun_out/r2a_bench_c3.log
tail -5 gpurun_out/r2a_pytest.log; tail -3 gpurun_out/r2a_ref_golden.log; tail -3 gpurun_out/r2a_ref_time.log; cat gpurun_out/r2a_mem_probe.log | head -12; cut -c1-600 gpurun_out/r2a_bench_c3.log
