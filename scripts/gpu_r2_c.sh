#!/bin/bash
# round 2, call C: whole GPU suite, bench lines (default C3, opaque C3, C5), reference arm through the reference's CUDA build.
(timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -60) > gpurun_out/r2c_pytest.log
(timeout 400 python bench.py 2>&1 | tail -3) > gpurun_out/r2c_bench_c3.json
(timeout 400 python bench.py --scene opaque --no-cpu 2>&1 | tail -3) > gpurun_out/r2c_bench_c3_opaque.json
(timeout 400 python bench.py --workload C5 --no-cpu 2>&1 | tail -3) > gpurun_out/r2c_bench_c5.json
(timeout 600 python bench.py --impl reference --steps 16 --warmup 4 2>&1 | tail -3) > gpurun_out/r2c_bench_reference.json
tail -15 gpurun_out/r2c_pytest.log; for f in gpurun_out/r2c_bench_*.json; do echo $f; cut -c1-700 $f; done
