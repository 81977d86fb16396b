#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== tests"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
echo "== poses"; python scripts/pose_samples.py C3 | python -c "
import json,sys
for i,p in enumerate(json.load(sys.stdin)): print(i,p['samples'],round(p['ms_total'],2),round(p['ms_setup'],2),round(p['ms_march'],2),round(p['ms_network'],2))"
echo "== bench C3"; timeout 900 python bench.py --steps 8 --warmup 3 --no-cpu > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo rc=$?; cut -c1-300 gpurun_out/bench_c3.json; tail -5 gpurun_out/bench_c3.err
