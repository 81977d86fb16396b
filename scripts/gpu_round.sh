#!/bin/bash
# One GPU-box call: smoke + parity tests + bench (C3 default, C2) + ncu launch list. Outputs under gpurun_out/.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader
echo "== smoke"; timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
echo "== tests"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
echo "== bench C3"; timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo rc=$?; cat gpurun_out/bench_c3.json; tail -5 gpurun_out/bench_c3.err
echo "== bench C2"; timeout 600 python bench.py --steps 8 --warmup 3 --workload C2 --no-cpu > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo rc=$?; cat gpurun_out/bench_c2.json; tail -5 gpurun_out/bench_c2.err
echo "== ncu launches"; timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu.log 2>&1; echo ncu rc=$?; tail -3 gpurun_out/ncu.log
