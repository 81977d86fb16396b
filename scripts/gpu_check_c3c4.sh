#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== tests"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
for w in C3 C4; do
echo "== bench $w"; timeout 900 python bench.py --workload $w --steps 32 --warmup 4 --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo rc=$?; python - $w <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/bench_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d.get('overflowed_timed_frames'))
PY
tail -3 gpurun_out/bench_$w.err
done
