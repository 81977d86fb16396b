#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== tests"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
echo "== mem probe"; timeout 300 python scripts/mem_probe.py 2>&1 | tee gpurun_out/mem_probe.log | tail -8
echo "== bench C3"; timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo rc=$?; cut -c1-300 gpurun_out/bench_c3.json; tail -5 gpurun_out/bench_c3.err
