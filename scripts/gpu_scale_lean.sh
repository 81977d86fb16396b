#!/bin/bash
# N-GPU box: the default bench line at N = 1, 2, 4, 8 exactly as the driver launches it (no tests, no CPU leg).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for n in ${NLIST:-1 2 4 8}; do
  if [ $n -eq 1 ]; then timeout 300 python bench.py --steps 16 --warmup 4 --no-cpu > gpurun_out/scale8_n$n.json 2> gpurun_out/scale8_n$n.err
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520+n)) bench.py --gpus $n --steps 16 --warmup 4 --no-cpu > gpurun_out/scale8_n$n.json 2> gpurun_out/scale8_n$n.err; fi
  echo "n=$n rc=$?"; python -c "
import json
d=json.loads(open('gpurun_out/scale8_n$n.json').read().strip().splitlines()[-1])
print('n',d['n_gpus'],'ms/frame',round(d['ms_per_step'],3),'Mrays/s',round(d['value']/1e6,1),'e2e',round(d['e2e']['value']/1e6,1))" 2>&1 | tail -1
done
