#!/bin/bash
# round 2, call R (N GPUs, default 2): per-device enqueue threads of the in-process mode: multi-device tests, inproc line with
# and without the threads (NRF_NO_ENQUEUE_THREADS=1).
N=${1:-2}
mkdir -p gpurun_out
[ "$N" = 2 ] && (timeout 900 python -m pytest tests -m gpu -q -k "two_gpus or batch or server" 2>&1 | tail -4) > gpurun_out/r2r_pytest_n$N.log
(timeout 400 python bench.py --gpus $N --mode inproc 2>&1 | tail -1) > gpurun_out/r2r_inproc_n$N.json
(NRF_NO_ENQUEUE_THREADS=1 timeout 400 python bench.py --gpus $N --mode inproc 2>&1 | tail -1) > gpurun_out/r2r_inproc_nothreads_n$N.json
cat gpurun_out/r2r_pytest_n$N.log 2>/dev/null
python - <<PY
import json
for f in ('r2r_inproc_n$N','r2r_inproc_nothreads_n$N'):
    d=json.loads(open(f'gpurun_out/{f}.json').read().strip().splitlines()[-1])
    print(f, round(d['ms_per_step'],3), round(d['value']/1e6,1), 'e2e', round(d['e2e']['value']/1e6,1), 'batch', round(d['e2e']['batch_call_value']/1e6,1))
PY
