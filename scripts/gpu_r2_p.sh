#!/bin/bash
# round 2, call P (2 GPUs): the batch pipeline with its collector thread: batch / server / two-device tests, inproc N=2 line,
# single-GPU default line (batch_call_value beside e2e).
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -k "batch or two_gpus or server or exercise" 2>&1 | tail -5) > gpurun_out/r2p_pytest.log
(timeout 400 python bench.py --gpus 2 --mode inproc 2>&1 | tail -1) > gpurun_out/r2p_n2_inproc.json
(timeout 400 python bench.py --no-cpu 2>&1 | tail -1) > gpurun_out/r2p_n1.json
cat gpurun_out/r2p_pytest.log
python - <<'PY'
import json
for f in ('r2p_n2_inproc','r2p_n1'):
    d=json.loads(open(f'gpurun_out/{f}.json').read().strip().splitlines()[-1])
    print(f, round(d['ms_per_step'],3), round(d['value']/1e6,1), 'e2e', round(d['e2e']['value']/1e6,1), 'batch', round(d['e2e']['batch_call_value']/1e6,1))
PY
