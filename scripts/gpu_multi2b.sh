#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
for g in peer; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 16 --warmup 4 --gather $g > gpurun_out/scale2_$g.json 2> gpurun_out/scale2_$g.err; echo "$g rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/scale2_$g.json').read().strip().splitlines()[-1])
print('$g n',d['n_gpus'],'ms/frame',round(d['ms_per_step'],3),'Mrays/s',round(d['value']/1e6,1),'e2e',round(d['e2e']['value']/1e6,1), d['frame_breakdown_ms'])"
tail -3 gpurun_out/scale2_$g.err | cut -c1-300
done
