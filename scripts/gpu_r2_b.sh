#!/bin/bash
# round 2, call B: goldens from the reference CUDA build (with device level scales + rays_d), parity tests under the
# reference-CUDA numerics (CPU half of the golden test included), short C3 bench.
mkdir -p gpurun_out/ref_cuda
(timeout 300 python baseline/run_ref_cuda.py golden --out gpurun_out/ref_cuda 2>&1 | tail -20) > gpurun_out/r2b_ref_golden.log
cp gpurun_out/ref_cuda/ref_cuda_*.npz tests/golden/
(timeout 600 python -m pytest tests/test_ref_cuda_golden.py -q -s 2>&1 | tail -60) > gpurun_out/r2b_golden_tests.log
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40) > gpurun_out/r2b_pytest.log
(timeout 300 python bench.py --steps 16 --warmup 4 --no-cpu 2>&1 | tail -5) > gpurun_out/r2b_bench_c3.log
tail -3 gpurun_out/r2b_ref_golden.log; tail -30 gpurun_out/r2b_golden_tests.log; tail -12 gpurun_out/r2b_pytest.log; cut -c1-400 gpurun_out/r2b_bench_c3.log
