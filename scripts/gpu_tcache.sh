#!/bin/bash
# A/B of the sample-t cache between the two march passes: parity tests, per-pose stage times and the C3 bench, with and
# without the cache (NRF_NO_TCACHE=1 re-marches in pass 2).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== tests"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for mode in on off; do
  if [ $mode = off ]; then export NRF_NO_TCACHE=1; else unset NRF_NO_TCACHE; fi
  echo "== poses tcache $mode"; python scripts/pose_samples.py C3 | python -c "
import json,sys
for i,p in enumerate(json.load(sys.stdin)): print(i,p['samples'],round(p['ms_total'],2),round(p['ms_setup'],2),round(p['ms_march'],2),round(p['ms_network'],2))"
  echo "== bench C3 tcache $mode"; timeout 900 python bench.py --steps 8 --warmup 3 --no-cpu > gpurun_out/bench_tc_$mode.json 2> gpurun_out/bench_tc_$mode.err; echo rc=$?; cut -c1-200 gpurun_out/bench_tc_$mode.json; tail -3 gpurun_out/bench_tc_$mode.err
done
