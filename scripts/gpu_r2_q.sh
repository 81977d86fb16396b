#!/bin/bash
# round 2, call Q (8 GPUs): final scaling lines with the layout-specialised k_network: ranks N = 1, 2, 4, 8 and inproc N = 8.
mkdir -p gpurun_out
TR() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n "$@"; }
(timeout 300 python bench.py --no-cpu 2>&1 | tail -1) > gpurun_out/r2q_scale_n1.json
for n in 2 4 8; do (TR $n bench.py --gpus $n 2>&1 | tail -1) > gpurun_out/r2q_scale_n$n.json; done
(timeout 300 python bench.py --gpus 8 --mode inproc 2>&1 | tail -1) > gpurun_out/r2q_scale_n8_inproc.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2q_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r2q_')[1], d.get('n_gpus'), round(d['ms_per_step'],3), round(d['value']/1e6,1), 'e2e', round(d['e2e']['value']/1e6,1), d['e2e'].get('batch_call_value'))
    except Exception as e: print(f, 'ERR', e, open(f).read()[-300:])
PY
