#!/bin/bash
# round 2, call M: what the driver runs at round end — GPU suite, smoke, reference arm, default bench.
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -8) > gpurun_out/r2m_pytest.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2) > gpurun_out/r2m_smoke.log
(timeout 900 python bench.py --impl reference --gpus 1 --steps 16 --warmup 4 2>&1 | tail -1) > gpurun_out/r2m_bench_reference.json
(timeout 600 python bench.py --gpus 1 --steps 16 --warmup 4 2>&1 | tail -1) > gpurun_out/r2m_bench_c3.json
tail -4 gpurun_out/r2m_pytest.log; cat gpurun_out/r2m_smoke.log; cut -c1-330 gpurun_out/r2m_bench_reference.json; echo; cut -c1-330 gpurun_out/r2m_bench_c3.json
